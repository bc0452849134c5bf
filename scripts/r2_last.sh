#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q --timeout=200 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -10; grep -E "^E  " gpurun_out/pytest_gpu.log | head -20
timeout 100 python bench.py --steps 500 --no-modes --no-eager --no-cpu-baseline 2>/dev/null | python -c "import json,sys;j=json.loads(sys.stdin.read());print(j['value'], j['ms_per_step'], j['e2e']['value'])"
