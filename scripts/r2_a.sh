#!/bin/bash
# Round-2 first GPU call: GPU tests (all failures listed), smoke, bench line, A/B of the new switches, timeline.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest -m gpu"
timeout 1500 python -m pytest tests -m gpu -q --timeout=400 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -40 gpurun_out/pytest_gpu.log
echo "== smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -6 gpurun_out/smoke.log
echo "== bench"
timeout 900 python bench.py > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?"
tail -5 gpurun_out/bench_a.err; cut -c1-1500 gpurun_out/bench_a.json
for v in "MDT_NO_PRELOAD=1" "MDT_STEPS_PER_GRAPH=1" "MDT_STEPS_PER_GRAPH=2" "MDT_STEPS_PER_GRAPH=32"; do
  echo "== $v"
  env $v timeout 300 python bench.py --steps 1000 --no-modes --no-cpu-baseline --no-eager > gpurun_out/bench_ab.json 2> gpurun_out/bench_ab.err; echo "rc=$?"
  python -c "import json;j=json.load(open('gpurun_out/bench_ab.json'));print(j['value'], j['ms_per_step'], j['timing']['one_call_per_step_ms'], j['e2e']['value'], j['e2e']['sync_value'])" || tail -5 gpurun_out/bench_ab.err
done
echo "== timeline"
GRAPH=1 timeout 200 python scripts/tc05_phases.py tf32x3 > gpurun_out/timeline_pre.txt 2>&1; cat gpurun_out/timeline_pre.txt | cut -c1-260
MDT_NO_PRELOAD=1 GRAPH=1 timeout 200 python scripts/tc05_phases.py tf32x3 > gpurun_out/timeline_nopre.txt 2>&1; cat gpurun_out/timeline_nopre.txt | cut -c1-260
