#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -40
grep -E "^E  " gpurun_out/pytest_gpu.log | head -40
echo "== bench"
timeout 900 python bench.py > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_g.err | cut -c1-300
python -c "
import json;j=json.load(open('gpurun_out/bench_g.json'))
for k in ('value','ms_per_step','e2e','compat_mode','unchanged_script'): print(k, json.dumps(j.get(k))[:1200])
"
