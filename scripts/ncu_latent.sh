#!/bin/bash
# One `ncu --set full` capture of the latent kernels (source-level stall sampling).  Run alone in a gpurun call.
set -x
mkdir -p gpurun_out
MDT_LATENT_SC=${SC:-2} GRAPH=0 timeout 300 python scripts/tc05_phases.py tf32x3 > gpurun_out/pre_ncu.log 2>&1 || exit 1
MDT_LATENT_SC=${SC:-2} GRAPH=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:latent_ -s 6 -c 2 \
  -o gpurun_out/latent -f python scripts/tc05_phases.py tf32x3 > gpurun_out/ncu_latent.log 2>&1
tail -5 gpurun_out/ncu_latent.log
