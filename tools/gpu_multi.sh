#!/bin/bash
# 8-GPU box: 1-GPU kNN sweep, 8-GPU sharded kNN sweep (config 5), then the bench at 2 / 4 / 8 GPUs as the driver launches it.
mkdir -p gpurun_out
timeout 300 python tools/knn_sweep.py 2000 8000 32000 100000 200000 > gpurun_out/knn_sweep_1gpu_v8.jsonl 2> gpurun_out/knn_sweep_1gpu_v8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/knn_sweep.py 2000 8000 32000 100000 200000 > gpurun_out/knn_sweep_8gpu_v8.jsonl 2> gpurun_out/knn_sweep_8gpu_v8.err
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_${n}gpu_v8.json 2> gpurun_out/bench_${n}gpu_v8.err
  tail -c 600 gpurun_out/bench_${n}gpu_v8.json | head -c 300; echo
done
timeout 900 python tools/cpu_baselines.py > gpurun_out/cpu_baselines_v8.jsonl 2> gpurun_out/cpu_baselines_v8.err; cat gpurun_out/cpu_baselines_v8.jsonl | cut -c1-200
cat gpurun_out/knn_sweep_1gpu_v8.jsonl gpurun_out/knn_sweep_8gpu_v8.jsonl | cut -c1-140
