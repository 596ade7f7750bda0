#!/bin/bash
# Option sweep of the bench workload (run under gpurun): prints one compact line per setting.
# usage: scripts/sweep.sh "variant=2" "fused_strip" 16 32 64 ...
set -u
FIXED=$1; NAME=$2; shift 2
for v in "$@"; do
  OPTS=""
  for kv in $FIXED; do OPTS="$OPTS --option $kv"; done
  python bench.py --steps 2 --warmup 1 --nt 400 --no-cpu-baseline $OPTS --option $NAME=$v 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
r=d['roofline']
print('$FIXED $NAME=$v', 'Gupd/s=%.2f'%(d['value']/1e9), 'e2e=%.2f'%(d['e2e']['value']/1e9), 'ms/launch', {k:round(v,4) for k,v in r['all_ms_per_launch'].items()}, 'frac=%.3f'%r['frac'])
"
done
