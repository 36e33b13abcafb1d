set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sufficient or suffix_array or nonelastic or golden or cfg2 or random_small" > gpurun_out/c4_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c4_pytest.log
tail -5 gpurun_out/c4_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/c4_bench.json 2> gpurun_out/c4_bench.err; echo "bench exit $?"
cat gpurun_out/c4_bench.json | head -c 400; tail -5 gpurun_out/c4_bench.err
