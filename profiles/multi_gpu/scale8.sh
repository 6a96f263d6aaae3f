R="python -m torch.distributed.run --nnodes=1 --nproc-per-node"
N=$1
timeout 600 $R $N --master-addr 127.0.0.1 --master-port 29911 bench.py --gpus $N --steps 300 --warmup 5 > gpurun_out/bench_v9_${N}gpu.json 2> gpurun_out/bench_v9_${N}gpu.err
tail -3 gpurun_out/bench_v9_${N}gpu.err
LGK_DP_FUSED=0 timeout 300 $R $N --master-addr 127.0.0.1 --master-port 29912 bench.py --gpus $N --steps 300 --warmup 5 --no-extras > gpurun_out/bench_v9_${N}gpu_nccl.json 2> gpurun_out/bench_v9_${N}gpu_nccl.err
tail -2 gpurun_out/bench_v9_${N}gpu_nccl.err
