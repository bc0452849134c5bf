#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -20
grep -E "^E  " gpurun_out/pytest_gpu.log | head -30
echo "== bench"
timeout 900 python bench.py > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_f.err | cut -c1-300
python -c "
import json;j=json.load(open('gpurun_out/bench_f.json'))
for k in ('value','ms_per_step','timing','modes','trials_per_gpu','compat_mode','torch_cuda_eager','unchanged_script','roofline'): print(k, json.dumps(j.get(k))[:900])
print(j['e2e']['value'], j['e2e']['sync_value'])
for k in j['kernels']: print(k)
"
echo "== errors"
timeout 600 python scripts/measure_errors.py > gpurun_out/measured_errors.txt 2>&1; cat gpurun_out/measured_errors.txt
