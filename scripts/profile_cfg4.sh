#!/bin/bash
# cfg4 (lens in the loop): plain run, then one full ncu capture of the generic Jacobian kernel.
set -u
TAG=${1:-r1m}
CMD="python scripts/cfg4_profile.py 1000 500 3"
$CMD > gpurun_out/${TAG}_cfg4_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_jacobian_lens -s 3 -c 1 \
    -o gpurun_out/${TAG}_cfg4_k_jacobian_lens -f $CMD > gpurun_out/${TAG}_cfg4_ncu.log 2>&1
ncu -i gpurun_out/${TAG}_cfg4_k_jacobian_lens.ncu-rep --page details > gpurun_out/${TAG}_cfg4_k_jacobian_lens_ncu_details.txt 2>&1
tail -3 gpurun_out/${TAG}_cfg4_plain.log
