#!/bin/bash
# round-2 ncu captures of the assembly kernels of a cfg4 solve (host-driven loop so that ncu sees plain launches)
export MMB_LM_GRAPH=0
CMD="python scripts/cfg4_profile.py 1000 500 4"
$CMD > gpurun_out/r2d_plain.log 2>&1 || exit 1
for PAIR in k_colmax:k_colmax 'k_assemble:^k_assemble$' k_assemble_shared:k_assemble_shared; do
  K=${PAIR%%:*}; RX=${PAIR#*:}
  ncu --set full --clock-control none --import-source on -k regex:$RX -s 1 -c 1 \
      -o gpurun_out/r2d_$K -f $CMD > gpurun_out/r2d_ncu_$K.log 2>&1
  ncu -i gpurun_out/r2d_$K.ncu-rep --page details > gpurun_out/r2d_cfg4_${K}_ncu_details.txt 2>&1
  ncu -i gpurun_out/r2d_$K.ncu-rep --page source --csv > gpurun_out/r2d_cfg4_${K}_source.csv 2>/dev/null
  rm -f gpurun_out/r2d_$K.ncu-rep
  grep -n "Duration\|Compute (SM) Throughput\|Memory Throughput\|Issue Slots Busy\|Executed Ipc Active\|Achieved Occupancy\|Registers Per\|DRAM Throughput\|L2 Cache Throughput\|Stall\|Est. Speedup" gpurun_out/r2d_cfg4_${K}_ncu_details.txt | head -24
done
