#!/bin/bash
# Scaling visit on an 8-GPU box: bench at N=4 and N=8 (dense512), then BASELINE config 5 (wdas_cloud_4k) at N=8.
mkdir -p gpurun_out
for N in 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "N=$N rc=$?"
  cut -c1-200 gpurun_out/bench_n$N.json
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus 8 --config wdas_cloud_4k --waves-per-step 64 --steps 3 --warmup 3 > gpurun_out/bench_c5_n8.json 2> gpurun_out/bench_c5_n8.err; echo "C5 N=8 rc=$?"
cut -c1-200 gpurun_out/bench_c5_n8.json
