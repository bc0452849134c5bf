#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29531 scripts/nccl_h4096_check.py 2>&1 | grep -E "nccl_h4096_check|Error|error" | head -8
echo "== pytest multi (<= 4 GPUs)"
timeout 500 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=280 -p no:cacheprovider -k "not eight" > gpurun_out/pytest_4gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_4gpu.log | head -20
grep -E "^E  |--- rank|Error" gpurun_out/pytest_4gpu.log | head -40
echo "== bench N=4"
timeout 420 $TR --master-port 29532 bench.py --gpus 4 --steps 300 --warmup 20 --max-seconds 380 > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "bench rc=$?"
grep -E "Traceback|Error|error" gpurun_out/bench_n4.err | grep -v NCCL | head -10
python -c "
import json;j=json.load(open('gpurun_out/bench_n4.json'))
for k in ('value','ms_per_step','incomplete','ddp_parity','layouts','layouts_bf16','exchanges','allreduce','torch_cuda_eager','cpu_baseline'): print(k, json.dumps(j.get(k))[:2500])
print(j['e2e']['value'], j['e2e']['sync_value'])
"
