#!/bin/bash
# per-kernel device times (cold cache, serialised): compare SHARES.  usage: ncu_list.sh <gemm-mode> [extra bench args]
m=${1:-tf32x3}; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 20 --warmup 5 --gemm-mode $m --no-cpu-baseline $@"
$CMD > gpurun_out/plain_$m.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_$m.csv $CMD > gpurun_out/ncu_$m.log 2>&1
echo ncu_rc=$?
python scripts/ncu_summary.py gpurun_out/launches_$m.csv
