# config 3 (GIN / SAGE, 4096 graphs per step) on N GPUs, weak and strong; usage: bash profiles/multi_gpu/cfg3_scale.sh N
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node"
N=$1
port=29930
for wl in gin_b4096 sage_b4096; do
  for sc in weak strong; do
    port=$((port+1))
    timeout 300 $R $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --workload $wl --scaling $sc --steps 60 --warmup 5 --no-extras --cpu-seconds 0.5 > gpurun_out/cfg3_${wl}_n${N}_${sc}.json 2> gpurun_out/cfg3_${wl}_n${N}_${sc}.err
    tail -1 gpurun_out/cfg3_${wl}_n${N}_${sc}.err | cut -c1-200
  done
done
