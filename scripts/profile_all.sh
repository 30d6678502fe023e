#!/bin/bash
# Run on the GPU box (under gpurun): everything profiles/ summarises for one round tag.
#   scripts/profile_all.sh <tag> [nobench]
# Writes into gpurun_out/: bench_<tag>.json (plain run, the only source of bench values), launches_<tag>.csv (ncu launch
# list of the same command, cold-cache and serialised: compare shares), prof_<tag>_<kernel>.ncu-rep + summaries, and
# traffic_<tag>.json (copy to profiles/traffic.json: bench.py's roofline blocks read it).
set -u
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
if [ "${2:-}" != "nobench" ]; then
  python bench.py --steps 10 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
  echo "bench rc=$?"; tail -c 1500 $OUT/bench_$TAG.err
fi
if [ "${NOLIST:-}" = "" ]; then
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extra"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
fi
M=smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__sass_thread_inst_executed_op_dadd_pred_on.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__sass_inst_executed_op_local_ld.sum,smsp__sass_inst_executed_op_local_st.sum
rm -f $OUT/traffic_$TAG.json
for T in ${TARGETS:-fixed adaptive_benign adaptive_bounds sim_angle}; do
  K='like_batch|like_lean'; KEY=like_batch_$T
  if [ $T = sim_angle ]; then K=sim_kernel; KEY=sim_kernel_angle; fi
  UNITS=$(python scripts/prof_target.py $T 2> $OUT/plain_${TAG}_$T.log | tail -1)
  echo "$T units/launch: $UNITS"
  ncu --set full --metrics $M --clock-control none --import-source on -k "regex:$K" -s 2 -c 1 -o $OUT/prof_${TAG}_$T -f python scripts/prof_target.py $T > $OUT/ncu_full_${TAG}_$T.log 2>&1
  echo "$T capture rc=$?"
  python scripts/ncu_summary.py $OUT/prof_${TAG}_$T.ncu-rep > $OUT/ncu_summary_${TAG}_$T.txt 2>&1
  python scripts/ncu_facts.py $OUT/traffic_$TAG.json $KEY $OUT/prof_${TAG}_$T.ncu-rep $UNITS
done
if [ -f $OUT/prof_${TAG}_fixed.ncu-rep ]; then
python scripts/line_profile.py $OUT/prof_${TAG}_fixed.ncu-rep 655360000 like_lean like_lean_kernel > $OUT/line_profile_${TAG}_fixed.txt 2>&1
python scripts/opcode_profile.py $OUT/prof_${TAG}_fixed.ncu-rep 655360000 > $OUT/opcode_profile_${TAG}_fixed.txt 2>&1
fi
ls -la $OUT | tail -15
