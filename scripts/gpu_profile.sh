#!/bin/bash
# Round profile recipe (run under gpurun): plain run first, then the ncu launch list and the
# full-set captures of the sweep kernels on the same command line.
set -u
mkdir -p gpurun_out
TAG=${1:-r1}
CMD="python bench.py --steps 1 --warmup 1 --nt 200 --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.log 2> gpurun_out/plain_$TAG.err || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 1000 -c 1500 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:march -s 400 -c 6 \
    -f -o gpurun_out/prof_fwd_$TAG $CMD > gpurun_out/ncu_fwd_$TAG.log 2>&1
echo "fwd capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:stress_march -s 419 -c 2 \
    -f -o gpurun_out/prof_adj_$TAG $CMD > gpurun_out/ncu_adj_$TAG.log 2>&1
echo "adj capture rc=$?"
ls -la gpurun_out
