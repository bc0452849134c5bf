#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_steps.py -m gpu -q --timeout=300 -p no:cacheprovider -x > gpurun_out/pytest_steps.log 2>&1; echo "steps rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_steps.log | head; grep -E "^E  " gpurun_out/pytest_steps.log | head -30
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -20
echo "== bench"
timeout 900 python bench.py --no-eager --no-cpu-baseline > gpurun_out/bench_h.json 2> gpurun_out/bench_h.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_h.err | cut -c1-300
python -c "
import json;j=json.load(open('gpurun_out/bench_h.json'))
for k in ('value','ms_per_step','trials_per_gpu','trials_per_gpu_grouped','modes'): print(k, json.dumps(j.get(k))[:600])
print(j['e2e']['value'])
"
