#!/bin/bash
mkdir -p gpurun_out
echo "== pytest -m gpu"
timeout 1800 python -m pytest tests -m gpu -q --timeout=400 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -60
grep -E "^E  " gpurun_out/pytest_gpu.log | head -80
echo "== bench"
timeout 900 python bench.py --no-eager > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?"
tail -5 gpurun_out/bench_b.err
python -c "
import json;j=json.load(open('gpurun_out/bench_b.json'))
for k in ('value','ms_per_step','modes','trials_per_gpu','compat_mode'): print(k, j[k])
print(j['e2e']['value'], j['e2e']['sync_value'])
for k in j['kernels']: print(k)
"
echo "== bench bf16"
timeout 600 python bench.py --gemm-mode bf16 --no-eager --no-modes --no-cpu-baseline > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "rc=$?"
tail -3 gpurun_out/bench_bf16.err
python -c "
import json;j=json.load(open('gpurun_out/bench_bf16.json'))
for k in ('value','ms_per_step'): print(k, j[k])
print(j['e2e']['value'], j['e2e']['sync_value'])
for k in j['kernels']: print(k)
"
GRAPH=1 timeout 200 python scripts/tc05_phases.py bf16 > gpurun_out/timeline_bf16.txt 2>&1; cut -c1-260 gpurun_out/timeline_bf16.txt
