set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "writer or cli_device or golden" > gpurun_out/c2_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2_pytest.log
tail -4 gpurun_out/c2_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err; echo "bench exit $?"
cat gpurun_out/c2_bench.json; tail -5 gpurun_out/c2_bench.err
(python tools/cli_e2e.py cfg2; python tools/cli_e2e.py cfg3 0.2) > gpurun_out/c2_cli.log 2>&1
cat gpurun_out/c2_cli.log
