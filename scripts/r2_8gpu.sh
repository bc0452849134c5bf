#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l
echo "== bench N=8"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps ${STEPS:-400} --warmup 20 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench rc=$?"
grep -v "^\[rank\|NCCL\|^$\|\*\*\*\|OMP_NUM" gpurun_out/bench_n8.err | tail -15 | cut -c1-300
python -c "
import json;j=json.load(open('gpurun_out/bench_n8.json'))
for k in ('value','ms_per_step','ddp_parity','layouts','layouts_bf16','exchanges','allreduce','config5','torch_cuda_eager','cpu_baseline'): print(k, json.dumps(j.get(k))[:3000])
print(j['e2e']['value'], j['e2e']['sync_value'])
"
echo "== pytest 4/8 GPU"
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=600 -p no:cacheprovider -k "four or eight or two_groups" > gpurun_out/pytest_8gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_8gpu.log | head -20
grep -E "^E  " gpurun_out/pytest_8gpu.log | head -30
