#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -10; grep -E "^E  " gpurun_out/pytest_gpu.log | head -20
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 400 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_final.err | cut -c1-300
python -c "
import json;j=json.load(open('gpurun_out/bench_final.json'))
for k in ('value','ms_per_step','modes','trials_per_gpu','trials_per_gpu_grouped','trials_per_gpu_streams','compat_mode','torch_cuda_eager','unchanged_script','roofline','cpu_baseline'): print(k, json.dumps(j.get(k))[:700])
print(j['e2e']['value'], j['e2e']['one_step_ahead_value'], j['e2e']['sync_value'], j['gpu_launches_per_step'])
"
