#!/bin/bash
# round-2 ncu captures of the rows kernel with all cameras in one launch: dense (mc) and sparse visibility (mcs)
for W in mc mcs; do
  python scripts/profile_r2.py $W > gpurun_out/r2c_${W}_plain.log 2>&1 || exit 1
  ncu --set full --clock-control none --import-source on -k regex:k_jacobian_rows -s 2 -c 1 \
      -o gpurun_out/r2c_${W}_rows python scripts/profile_r2.py $W > gpurun_out/r2c_${W}_ncu.log 2>&1
  ncu -i gpurun_out/r2c_${W}_rows.ncu-rep --page details > gpurun_out/r2c_${W}_k_jacobian_rows_ncu_details.txt 2>/dev/null
  ncu -i gpurun_out/r2c_${W}_rows.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed.sum > gpurun_out/r2c_${W}_rows_raw.csv 2>/dev/null
  rm -f gpurun_out/r2c_${W}_rows.ncu-rep
done
cat gpurun_out/r2c_mc_plain.log gpurun_out/r2c_mcs_plain.log
cat gpurun_out/r2c_mc_rows_raw.csv gpurun_out/r2c_mcs_rows_raw.csv
grep -n "Duration\|Compute (SM) Throughput\|Issue Slots Busy\|Executed Ipc Active\|Achieved Occupancy\|Registers Per\|Bank\|Stall" gpurun_out/r2c_mcs_k_jacobian_rows_ncu_details.txt | head -30
