#!/bin/bash
# Run on the GPU box (under gpurun): bench line, ncu launch list, one --set full capture of the batch kernel.
# Usage: scripts/profile.sh <tag>
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps 5 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 3000 $OUT/bench_$TAG.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --sets 8192"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
CMD2="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra --sets 4096"
$CMD2 > $OUT/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:like_batch -s 3 -c 1 -o $OUT/prof_$TAG -f $CMD2 > $OUT/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
ls -la $OUT | tail -12
