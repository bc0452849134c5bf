#!/bin/bash
# 8-GPU box: multi-GPU parity, whole-box bench at N=8, 4, 2 with every group layout.
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=300 2>&1 | tail -3
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  timeout 400 $TR --nproc-per-node $n --master-port 2951$n bench.py --gpus $n --steps 3000 --warmup 20 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo bench$n rc=$?
  python -c "
import json; j=json.load(open('gpurun_out/bench_n$n.json')); print(j['value'], j['layouts'], round(j['e2e']['value']))"
done
