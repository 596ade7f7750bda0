#!/bin/bash
# Everything round 2 wants measured on the 8-GPU box, in one gpurun --gpus 8 call (charged 8x: keep it short).
OUT=gpurun_out/r2y
mkdir -p $OUT
cd "$(dirname "$0")/.."
nvidia-smi -L | wc -l > $OUT/gpus.txt
(timeout 300 python -m pytest tests/test_gpu_devices.py tests/test_gpu_slab.py -x -q > $OUT/pytest.log 2>&1; echo "rc=$?" >> $OUT/pytest.log)
tail -3 $OUT/pytest.log
timeout 300 python scripts/multi_device_bench.py 2000 1 2>/dev/null | grep '^{' > $OUT/one_object.jsonl
cat $OUT/one_object.jsonl
timeout 400 python scripts/slab_inprocess_bench.py --check --nt 1000 --counts 1,2,4,8 2>/dev/null | grep '^{' > $OUT/slab_inprocess.jsonl
timeout 200 python scripts/slab_inprocess_bench.py --nt 1000 --counts 8 --halo-steps 8 2>/dev/null | grep '^{' >> $OUT/slab_inprocess.jsonl

cut -c1-420 $OUT/slab_inprocess.jsonl
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 2 --warmup 3 > $OUT/bench_n8.json 2> $OUT/bench_n8.err
tail -1 $OUT/bench_n8.json | cut -c1-300
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2y/bench_n8.json").read().strip().splitlines()[-1])
    print("N=8 value", d["value"] / 1e9, "e2e", d["e2e"]["value"] / 1e9, json.dumps(d.get("extra"))[:900])
except Exception as e:
    print("bench n8 parse failed", e)
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > $OUT/bench_ref_n8.json 2> $OUT/bench_ref_n8.err
tail -1 $OUT/bench_ref_n8.json | cut -c1-900
