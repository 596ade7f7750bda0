#!/bin/bash
# compute-sanitizer over the default step kernel and its relatives (run on the GPU box):
#   scripts/sanitize.sh [out_dir]      -> <out_dir>/sanitizer_{memcheck,racecheck,synccheck,initcheck}.log
# memcheck covers everything; racecheck (shared-memory hazards: the TMA ring, the resident kernel's tiles) and
# synccheck (barriers) run the kernel families one by one to bound the run time.
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
cd "$(dirname "$0")/.."
export PSV_SANITIZE_NT=${PSV_SANITIZE_NT:-24}
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck synccheck racecheck; do
  log="$OUT/sanitizer_$tool.log"
  : > "$log"
  for what in pipe batch slab resident pair; do
    echo "==== $tool / $what ====" >> "$log"
    timeout 900 $CS --tool $tool --print-limit 20 python scripts/sanitize_case.py $what >> "$log" 2>&1
    echo "exit code: $?" >> "$log"
  done
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|exit code|^ok:|====" "$log" | tail -40
done
