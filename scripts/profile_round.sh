#!/bin/bash
# Round profile bundle (run under gpurun, 1 GPU): bench line, ncu launch list, ncu --set full of one step's GEMMs.
R=${1:-r1}
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench_$R.json 2> gpurun_out/bench_$R.err; echo bench_rc=$?; tail -2 gpurun_out/bench_$R.err
CMD="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-modes"
$CMD > gpurun_out/plain_$R.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_$R.csv $CMD > gpurun_out/ncu_list_$R.log 2>&1
echo list_rc=$?
python scripts/ncu_summary.py gpurun_out/launches_$R.csv > gpurun_out/launches_${R}_summary.txt; cat gpurun_out/launches_${R}_summary.txt
$CMD > gpurun_out/plain2_$R.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"tc05_gemm_kernel|latent_|prep_kernel" -s 100 -c 10 -o gpurun_out/prof_$R $CMD > gpurun_out/ncu_full_$R.log 2>&1
echo full_rc=$?; ls -la gpurun_out/prof_$R.ncu-rep
