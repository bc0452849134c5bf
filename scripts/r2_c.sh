#!/bin/bash
mkdir -p gpurun_out
echo "== pytest -m gpu"
timeout 1800 python -m pytest tests -m gpu -q --timeout=400 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -60
grep -E "^E  " gpurun_out/pytest_gpu.log | head -60
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
for m in tf32x3 bf16; do
for v in "X=1" "MDT_NO_PIPELINE=1" "MDT_DW3_LATE=1"; do
  echo "== $m $v"
  env $v timeout 300 python bench.py --gemm-mode $m --steps 1000 --no-modes --no-cpu-baseline --no-eager > gpurun_out/bench_c_$m.json 2> gpurun_out/bench_c.err; echo "rc=$?"
  python -c "import json;j=json.load(open('gpurun_out/bench_c_$m.json'));print(j['value'], j['ms_per_step'], j['timing']['one_call_per_step_ms'], j['e2e']['value'], j['e2e']['sync_value'], j['gpu_launches_per_step'])" || tail -5 gpurun_out/bench_c.err
done
GRAPH=1 timeout 200 python scripts/tc05_phases.py $m > gpurun_out/timeline_c_$m.txt 2>&1; cut -c1-200 gpurun_out/timeline_c_$m.txt
SOURCED=1 GRAPH=1 timeout 200 python scripts/tc05_phases.py $m > gpurun_out/timeline_c_src_$m.txt 2>&1; cut -c1-200 gpurun_out/timeline_c_src_$m.txt
done
