#!/bin/bash
# usage (under gpurun): bash tools/gpu_step_profile.sh <tag> [launches_per_step]
# plain run first (exit code checked), then the ncu launch list of the third step.
TAG=${1:-x}; LPS=${2:-72}; PAIRS=${PAIRS:-30}
python tools/profile_step.py 3 $PAIRS > gpurun_out/plain_$TAG.log 2>&1 || { cat gpurun_out/plain_$TAG.log; exit 1; }
LPS=$(awk '/launches_per_step/{print $2}' gpurun_out/plain_$TAG.log)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none -s $((2*LPS)) -c $LPS --csv --log-file gpurun_out/launches_$TAG.csv python tools/profile_step.py 3 $PAIRS > gpurun_out/ncu_$TAG.log 2>&1
python tools/ncu_launches.py gpurun_out/launches_$TAG.csv > gpurun_out/launches_$TAG.txt
tail -12 gpurun_out/launches_$TAG.txt
