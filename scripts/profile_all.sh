#!/bin/bash
# Run on the GPU box (under gpurun): everything profiles/ summarises for one round tag.
# Usage: scripts/profile_all.sh <tag>
set -u
TAG=${1:-r1}
bash scripts/profile.sh $TAG
bash scripts/profile_sim.sh $TAG
python scripts/dev_hostile_one.py > gpurun_out/plain_hostile_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:like_batch -s 2 -c 1 -o gpurun_out/prof_hostile_$TAG -f python scripts/dev_hostile_one.py > gpurun_out/ncu_hostile_$TAG.log 2>&1
echo "hostile capture rc=$?"
python scripts/dev_latency.py > gpurun_out/latency_$TAG.log 2>&1; python scripts/dev_latency3.py >> gpurun_out/latency_$TAG.log 2>&1; python scripts/dev_sweep.py >> gpurun_out/latency_$TAG.log 2>&1
python scripts/dev_hostile.py > gpurun_out/hostile_times_$TAG.log 2>&1
cat gpurun_out/latency_$TAG.log gpurun_out/hostile_times_$TAG.log
