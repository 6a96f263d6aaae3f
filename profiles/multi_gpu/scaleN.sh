R="python -m torch.distributed.run --nnodes=1 --nproc-per-node"
N=$1
timeout 400 $R $N --master-addr 127.0.0.1 --master-port 29921 bench.py --gpus $N --steps 300 --warmup 5 --no-extras --cpu-seconds 1 > gpurun_out/bench_v10_${N}gpu.json 2> gpurun_out/bench_v10_${N}gpu.err
tail -2 gpurun_out/bench_v10_${N}gpu.err
