set -x
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sufficient" > gpurun_out/c3_pytest_a.log 2>&1; echo "pytest exit $?" >> gpurun_out/c3_pytest_a.log
tail -30 gpurun_out/c3_pytest_a.log
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/c3_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c3_pytest.log
tail -15 gpurun_out/c3_pytest.log
timeout 120 python tools/fuzz_gpu.py 30 21000 > gpurun_out/c3_fuzz.log 2>&1; tail -12 gpurun_out/c3_fuzz.log
python bench.py --steps 5 --warmup 3 > gpurun_out/c3_bench.json 2> gpurun_out/c3_bench.err; echo "bench exit $?"
cat gpurun_out/c3_bench.json; tail -5 gpurun_out/c3_bench.err
