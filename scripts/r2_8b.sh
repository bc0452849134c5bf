#!/bin/bash
mkdir -p gpurun_out
timeout 215 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 300 --warmup 20 --no-eager --max-seconds 170 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench rc=$?"
grep -E "Traceback|Error|error|watchdog" gpurun_out/bench_n8.err | grep -v "NCCL INFO" | head -10
python -c "
import json;j=json.load(open('gpurun_out/bench_n8.json'))
for k in ('value','ms_per_step','incomplete','ddp_parity','layouts','layouts_bf16','exchanges','allreduce','config5','cpu_baseline'): print(k, json.dumps(j.get(k))[:3500])
print(j['e2e']['value'], j['e2e']['sync_value'])
"
