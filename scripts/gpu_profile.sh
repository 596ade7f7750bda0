#!/bin/bash
# Round profile recipe (run under gpurun): plain run first, then the ncu launch list and the
# full-set capture of the step kernels on the same command line.
#   scripts/gpu_profile.sh TAG [bench options...]
set -u
mkdir -p gpurun_out
TAG=${1:-r1}; shift
CMD="python bench.py --steps 1 --warmup 1 --nt 200 --no-cpu-baseline $*"
$CMD > gpurun_out/plain_$TAG.log 2> gpurun_out/plain_$TAG.err || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 800 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "launch list rc=$?"
# forward: plain + snapshot launches; adjoint: plain + accumulate launches
ncu --set full --clock-control none --import-source on -k regex:'fused|march' -s 215 -c 7 \
    -f -o gpurun_out/prof_fwd_$TAG $CMD > gpurun_out/ncu_fwd_$TAG.log 2>&1
echo "fwd capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'fused|march' -s 1017 -c 5 \
    -f -o gpurun_out/prof_adj_$TAG $CMD > gpurun_out/ncu_adj_$TAG.log 2>&1
echo "adj capture rc=$?"
# text exports (read here without a GPU); the adjoint report itself stays on the box if the two
# reports together would exceed gpurun's 64 MiB return limit
for w in fwd adj; do
  ncu -i gpurun_out/prof_${w}_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_${w}_$TAG.csv 2>/dev/null
  ncu -i gpurun_out/prof_${w}_$TAG.ncu-rep --page details --csv > gpurun_out/details_${w}_$TAG.csv 2>/dev/null
done
total=$(du -sm gpurun_out | cut -f1)
if [ "$total" -gt 56 ]; then rm -f gpurun_out/prof_adj_$TAG.ncu-rep; fi
ls -la gpurun_out | tail -12
