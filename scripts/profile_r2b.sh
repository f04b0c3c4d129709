#!/bin/bash
# round-2 ncu capture of the per-frame LM kernel (24 000 frames) and launch list of the bench
python scripts/perframe_profile.py > gpurun_out/r2_pf_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_perframe_lm -s 4 -c 1 \
    -o gpurun_out/r2_pf python scripts/perframe_profile.py > gpurun_out/r2_pf_ncu.log 2>&1
ncu -i gpurun_out/r2_pf.ncu-rep --page details > gpurun_out/r2_k_perframe_lm_ncu_details.txt 2>/dev/null
rm -f gpurun_out/r2_pf.ncu-rep
grep -n "Duration\|Compute (SM) Throughput\|Issue Slots Busy\|Executed Ipc Active\|Achieved Occupancy\|Registers Per\|FP64\|Local\|Grid Size" gpurun_out/r2_k_perframe_lm_ncu_details.txt | head -20
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r2k_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2k_launches_bench.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r2k_ncu1.log 2>&1
python scripts/launch_totals.py gpurun_out/r2k_launches_bench.csv 12
