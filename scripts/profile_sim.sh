#!/bin/bash
# ncu --set full capture of the simulator kernel (angle model, 256 sets x 50k particles, histogram mode)
TAG=${1:-r1}
CMD="python scripts/dev_sim_time.py 256 50000"
$CMD > gpurun_out/plain_sim_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sim_kernel -s 1 -c 1 -o gpurun_out/prof_sim_$TAG -f $CMD > gpurun_out/ncu_sim_$TAG.log 2>&1
echo "rc=$?"; cat gpurun_out/plain_sim_$TAG.log
