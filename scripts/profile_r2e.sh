#!/bin/bash
# round-2 ncu capture of the SYRK of a cfg3 solve (host-driven loop so that ncu sees plain launches)
export MMB_LM_GRAPH=0
CMD="python scripts/solve_profile.py 1000 500 3"
$CMD > gpurun_out/r2e_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_schur_syrk -s 2 -c 1 -o gpurun_out/r2e_syrk -f $CMD > gpurun_out/r2e_ncu.log 2>&1
ncu -i gpurun_out/r2e_syrk.ncu-rep --page details > gpurun_out/r2e_k_schur_syrk_ncu_details.txt 2>&1
ncu -i gpurun_out/r2e_syrk.ncu-rep --page source --csv > gpurun_out/r2e_k_schur_syrk_source.csv 2>/dev/null
rm -f gpurun_out/r2e_syrk.ncu-rep
grep -n "Duration\|Compute (SM) Throughput\|Issue Slots Busy\|Executed Ipc Active\|Achieved Occupancy\|Registers Per\|Est. Speedup\|Stall\|stall\|Eligible\|Pipe" gpurun_out/r2e_k_schur_syrk_ncu_details.txt | head -40
