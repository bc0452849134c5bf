#!/bin/bash
# Round-2 profile bundle (1 GPU): ncu launch lists and `--set full` captures of one step's kernels, tf32x3 and bf16.
mkdir -p gpurun_out
for m in tf32x3 bf16; do
  CMD="python bench.py --gemm-mode $m --steps 20 --warmup 5 --repeats 1 --no-cpu-baseline --no-modes --no-eager"
  timeout 300 $CMD > gpurun_out/plain_$m.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$m.log; continue; }
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches_$m.csv $CMD > gpurun_out/ncu_list_$m.log 2>&1
  echo "list_rc=$?"
  python scripts/ncu_summary.py gpurun_out/r2_launches_$m.csv > gpurun_out/r2_ncu_launch_list_summary_$m.txt; tail -14 gpurun_out/r2_ncu_launch_list_summary_$m.txt
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tc05_gemm_kernel|latent_|prep_kernel" -s 100 -c 10 -f -o gpurun_out/r2_prof_$m $CMD > gpurun_out/ncu_full_$m.log 2>&1
  echo "full_rc=$?"
  ncu -i gpurun_out/r2_prof_$m.ncu-rep --page raw --csv > gpurun_out/r2_prof_raw_$m.csv 2>/dev/null
  python scripts/ncu_full_summary.py gpurun_out/r2_prof_raw_$m.csv > gpurun_out/r2_ncu_full_summary_$m.txt; cut -c1-260 gpurun_out/r2_ncu_full_summary_$m.txt
  gzip -f gpurun_out/r2_launches_$m.csv
done
ls -la gpurun_out/*.ncu-rep
