#!/bin/bash
# Multi-GPU pass on one box: strong scaling of the 1024-object map batch over 1/2/4/8 GPUs and the weak-scaling
# bench line at 8 (and 2) GPUs.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
: > gpurun_out/map_scale.jsonl
for n in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540 + n)) \
    scripts/bench_map_scale.py 2> gpurun_out/map_scale_$n.err | tail -1 >> gpurun_out/map_scale.jsonl
done
for n in 2 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29560 + n)) \
    bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_gpus$n.json 2> gpurun_out/bench_gpus$n.err
done
cat gpurun_out/map_scale.jsonl
tail -c 600 gpurun_out/bench_gpus8.json
