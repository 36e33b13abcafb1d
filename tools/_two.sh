set -x
mkdir -p gpurun_out
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/t_bench2.json 2> gpurun_out/t_bench2.err; echo "bench2 exit $?"
tail -c 1500 gpurun_out/t_bench2.json; tail -3 gpurun_out/t_bench2.err
timeout 100 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/t_sharded.log 2>&1; echo "sharded exit $?"; tail -3 gpurun_out/t_sharded.log
