#!/bin/bash
# full captures of the dominant LM kernels of one cfg3 global solve (host-driven loop so
# that ncu sees plain launches; the graph replays the same kernels)
TAG=${1:-r1n}
export MMB_LM_GRAPH=0
CMD="python scripts/solve_profile.py 1000 500 3"
$CMD > gpurun_out/${TAG}_solve_plain2.log 2>&1 || exit 1
for PAIR in k_schur_syrk:k_schur_syrk k_chol_tiles:k_chol_tiles 'k_assemble:^k_assemble$' k_assemble_shared:k_assemble_shared \
            k_block_factor:k_block_factor k_backward_wave:k_backward_wave k_forward_wave:k_forward_wave \
            k_gemv_t_partial:k_gemv_t_partial; do
  K=${PAIR%%:*}; RX=${PAIR#*:}
  ncu --set full --clock-control none --import-source on -k regex:$RX -s 1 -c 1 \
      -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
  ncu -i gpurun_out/${TAG}_$K.ncu-rep --page details > gpurun_out/${TAG}_${K}_ncu_details.txt 2>&1
done
ls gpurun_out | grep ${TAG} | tail -20
