R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
port=29800
for wl in gat3x64h4_ddr_b256_k6 gin_b4096 sage_b4096; do
  timeout 300 python bench.py --workload $wl --steps 100 --warmup 5 --no-extras --cpu-seconds 2 > gpurun_out/dp_${wl}_n1.json 2> gpurun_out/dp_${wl}_n1.err
  for sc in weak strong; do
    port=$((port+1))
    timeout 400 $R $port bench.py --gpus 2 --workload $wl --scaling $sc --steps 100 --warmup 5 > gpurun_out/dp_${wl}_n2_${sc}.json 2> gpurun_out/dp_${wl}_n2_${sc}.err
    tail -2 gpurun_out/dp_${wl}_n2_${sc}.err
  done
done
ls -la gpurun_out/dp_*.json
