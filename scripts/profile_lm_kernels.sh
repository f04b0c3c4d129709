#!/bin/bash
# full captures of the dominant LM kernels of one cfg3 global solve
TAG=${1:-r1e}
CMD="python scripts/solve_profile.py 1000 500 3"
MMB_LM_PROFILE=0 $CMD > gpurun_out/${TAG}_solve_plain2.log 2>&1 || exit 1
for K in k_chol_panel k_chol_trailing k_schur_syrk k_assemble; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 \
      -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
done
ls -la gpurun_out | tail -8
