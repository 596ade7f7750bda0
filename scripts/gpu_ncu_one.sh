#!/bin/bash
# One full-set ncu capture of the step kernel (run under gpurun): scripts/gpu_ncu_one.sh TAG [bench options]
set -u
mkdir -p gpurun_out
TAG=$1; shift
CMD="python bench.py --steps 1 --warmup 1 --nt 100 --no-cpu-baseline $*"
$CMD > gpurun_out/plain_$TAG.log 2> gpurun_out/plain_$TAG.err || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'window|fused|march' -s 113 -c 2 \
    -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "capture rc=$?"
