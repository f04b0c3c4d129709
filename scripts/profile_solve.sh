#!/bin/bash
# launch list of one cfg3 global solve (shares of the LM kernels)
TAG=${1:-r1k}
CMD="python scripts/solve_profile.py 1000 500 6"
MMB_LM_PROFILE=0 $CMD > gpurun_out/${TAG}_solve_plain.log 2>&1 &&
# conditional-node graphs cannot be profiled per kernel: the launch list comes from the host loop
MMB_LM_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 700 --csv \
    --log-file gpurun_out/${TAG}_launches_solve.csv $CMD > gpurun_out/${TAG}_ncu_solve.log 2>&1
tail -3 gpurun_out/${TAG}_solve_plain.log
