#!/bin/bash
# Round-3 ncu recipe (run under gpurun, one GPU), for the default FP64 step kernel and the optional FP32 wavefield mode:
# plain run first (must exit 0 without ncu), then the launch list, then ONE full-set capture of 6 step launches each.
set -u
OUT=gpurun_out/r3p
mkdir -p $OUT
sha256sum psvwave_b200/libpsvwave_b200.so | cut -c1-16 > $OUT/library_sha16.txt
for MODE in f64 f32; do
  OPT=""; [ $MODE = f32 ] && OPT="fp32=1"
  CMD="python scripts/ncu_target.py $OPT"
  $CMD > $OUT/plain_$MODE.log 2> $OUT/plain_$MODE.err || { echo "plain run ($MODE) failed"; tail -5 $OUT/plain_$MODE.err; exit 1; }
  tail -1 $OUT/plain_$MODE.log
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file $OUT/launches_$MODE.csv $CMD > $OUT/ncu_list_$MODE.log 2>&1
  echo "launch list ($MODE) rc=$?"
  ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'fused(32)?_pipe' -c 6 \
      -f -o $OUT/prof_$MODE $CMD > $OUT/ncu_full_$MODE.log 2>&1
  echo "full capture ($MODE) rc=$?"
  ncu -i $OUT/prof_$MODE.ncu-rep --page raw --csv > $OUT/raw_$MODE.csv 2>/dev/null
  ncu -i $OUT/prof_$MODE.ncu-rep --page details --csv > $OUT/details_$MODE.csv 2>/dev/null
done
ls -la $OUT | tail -20
