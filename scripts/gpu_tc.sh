#!/bin/bash
# tcgen05 path check: phase timing, parity tests in tc modes, bench per mode.  Logs in gpurun_out/.
mkdir -p gpurun_out
timeout 120 python scripts/tc05_phases.py > gpurun_out/phases.log 2>&1; cat gpurun_out/phases.log
MDT_TEST_MODES=${MODES:-tf32x3,tf32} timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout=300 > gpurun_out/pytest_tc.log 2>&1; echo pytest_rc=$?
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_tc.log | head; grep -E "^E  " gpurun_out/pytest_tc.log | cut -c1-200 | head -10
for m in ${BENCH_MODES:-tf32x3}; do
  timeout 300 python bench.py --steps 5000 --warmup 20 --gemm-mode $m --no-cpu-baseline > gpurun_out/bench_$m.json 2> gpurun_out/bench_$m.err; echo bench_rc=$?; tail -2 gpurun_out/bench_$m.err
  python - <<PY
import json
j=json.load(open("gpurun_out/bench_$m.json"))
print("$m", "value", j["value"], "us/step", round(j["ms_per_step"]*1e3,2), "e2e", round(j["e2e"]["value"]), "launches/step", j["gpu_launches"]/j["steps"])
print("   ", " ".join(f'{k["kernel"]}={k["avg_us"]}' for k in j["kernels"]))
PY
done
