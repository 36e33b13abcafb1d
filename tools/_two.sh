set -x
mkdir -p gpurun_out
env | grep -i nccl
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/t_bench2.json 2> gpurun_out/t_bench2.err; echo "bench2 exit $?"
wc -l gpurun_out/t_bench2.json; head -c 200 gpurun_out/t_bench2.json
