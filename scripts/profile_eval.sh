#!/bin/bash
# One gpurun call: plain bench run, then the ncu launch list and a full capture of the
# Jacobian kernel of the same command (B200_PROFILING.md recipe).
set -u
TAG=${1:-r1k}
CMD="python bench.py --steps 5 --warmup 3 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches_bench.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_jacobian_rows -s 4 -c 2 \
    -o gpurun_out/${TAG}_k_jacobian_rows -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_variants -s 4 -c 1 \
    -o gpurun_out/${TAG}_k_variants -f $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
ls -la gpurun_out/
