#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=400 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -20
grep -E "^E  " gpurun_out/pytest_gpu.log | head -30
for m in tf32x3 bf16; do
for v in "MDT_GRAPH_DBG=1" "MDT_NO_GRAPH_REDUCE=1" "MDT_STEPS_PER_GRAPH=32"; do
  echo "== $m $v"
  env $v timeout 300 python bench.py --gemm-mode $m --steps 1000 --no-modes --no-cpu-baseline --no-eager > gpurun_out/bench_e_$m.json 2> gpurun_out/bench_e.err; echo "rc=$?"
  grep "redundant" gpurun_out/bench_e.err | sort | uniq -c | head -5
  python -c "import json;j=json.load(open('gpurun_out/bench_e_$m.json'));print(j['value'], j['ms_per_step'], j['timing']['one_call_per_step_ms'], j['e2e']['value'], j['e2e']['sync_value'], j['gpu_launches_per_step'])" || tail -5 gpurun_out/bench_e.err
done
SOURCED=1 GRAPH=1 timeout 200 python scripts/tc05_phases.py $m > gpurun_out/timeline_e_src_$m.txt 2>&1; cut -c1-130 gpurun_out/timeline_e_src_$m.txt
done
