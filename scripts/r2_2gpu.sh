#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_reference_script.py -m gpu -q --timeout=600 -p no:cacheprovider > gpurun_out/pytest_2gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_2gpu.log | head -20
grep -E "^E  " gpurun_out/pytest_2gpu.log | head -40
echo "== bench N=2"
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 500 --warmup 20 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
tail -15 gpurun_out/bench_n2.err | cut -c1-300
python -c "
import json;j=json.load(open('gpurun_out/bench_n2.json'))
for k in ('value','ms_per_step','ddp_parity','layouts','layouts_bf16','exchanges','allreduce','torch_cuda_eager','cpu_baseline'): print(k, json.dumps(j.get(k))[:1500])
print(j['e2e']['value'], j['e2e']['sync_value'])
"
echo "== reference arm N=2"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 20 --warmup 3 2> gpurun_out/ref_n2.err | cut -c1-600
