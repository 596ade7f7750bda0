#!/bin/bash
# Round-2 ncu recipe (run under gpurun, one GPU): plain run, launch list, full-set captures of the three launch kinds.
set -u
OUT=gpurun_out/r2p
mkdir -p $OUT
CMD="python scripts/ncu_target.py $*"
$CMD > $OUT/plain.log 2> $OUT/plain.err || { echo "plain run failed"; tail -5 $OUT/plain.err; exit 1; }
tail -1 $OUT/plain.log
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file $OUT/launches.csv $CMD > $OUT/ncu_list.log 2>&1
echo "launch list rc=$?"
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:fused_pipe -c 6 \
    -f -o $OUT/prof_fwd $CMD > $OUT/ncu_fwd.log 2>&1
echo "fwd capture rc=$?"
ncu --profile-from-start off --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'fused_pipe_kernel<psv::FastMath, *\(int\)2|fused_pipe_kernel<psv::FastMath, 2' -c 2 \
    -f -o $OUT/prof_corr $CMD > $OUT/ncu_corr.log 2>&1
echo "corr capture rc=$?"
for w in fwd corr; do
  ncu -i $OUT/prof_$w.ncu-rep --page raw --csv > $OUT/raw_$w.csv 2>/dev/null
  ncu -i $OUT/prof_$w.ncu-rep --page details --csv > $OUT/details_$w.csv 2>/dev/null
done
ls -la $OUT | tail -12
