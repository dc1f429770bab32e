#!/bin/bash
# usage (under gpurun): bash tools/gpu_farneback_profile.sh <tag>   -- plain run first, then the ncu launch list of the last call
TAG=${1:-x}
python tools/profile_farneback.py 3 > gpurun_out/fb_plain_$TAG.log 2>&1 || { cat gpurun_out/fb_plain_$TAG.log; exit 1; }
LPS=$(awk '/launches_per_call/{print $2}' gpurun_out/fb_plain_$TAG.log)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none -s $((2*LPS)) -c $LPS --csv --log-file gpurun_out/fb_launches_$TAG.csv python tools/profile_farneback.py 3 > gpurun_out/fb_ncu_$TAG.log 2>&1
python tools/ncu_launches.py gpurun_out/fb_launches_$TAG.csv > gpurun_out/fb_launches_$TAG.txt
cat gpurun_out/fb_plain_$TAG.log; tail -14 gpurun_out/fb_launches_$TAG.txt
