"""Monte-Carlo ensemble throughput (BASELINE.json configs[3]), run under gpurun:
C1 geometry (default ini, 250x150 grid, nt=8000), B models = background*(1 + 0.05*smoothed N(0,1)),
rng 2024, all forward-simulated together by psv_forward_batch (one launch per time step for the
ensemble).  Prints one JSON line per setting: models/s and grid-point updates/s, device time from
CUDA events on the leader's stream, next to the same models simulated one after the other.
usage: python scripts/mc_bench.py [B] [nt] [name=value ...]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200  # noqa: E402
from helpers import C1_INI, make_config  # noqa: E402


def smooth(a, n=6):
    for _ in range(n):
        a = (a + np.roll(a, 1, 0) + np.roll(a, -1, 0) + np.roll(a, 1, 1) + np.roll(a, -1, 1)) / 5.0
    return a


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    nt = int(sys.argv[2]) if len(sys.argv) > 2 else 8000
    opts = dict(kv.split("=") for kv in sys.argv[3:])
    ini = "/tmp/mc_c1.ini"
    make_config(ini, nt=nt)
    rng = np.random.default_rng(2024)
    models = []
    for _ in range(B):
        m = psvwave_b200.fdModel(ini)
        for k, v in opts.items():
            m._set_option(k, int(v))
        f = [1 + 0.05 * smooth(rng.standard_normal((m.nx, m.nz))) / 0.2 for _ in range(3)]
        m.set_parameter_fields(2000.0 * f[0], 800.0 * f[1], 1500.0 * f[2])
        models.append(m)
    lead = models[0]
    pts = lead.nx * lead.nz
    psvwave_b200.forward_batch(models, 0, False)  # warm-up
    t0 = time.perf_counter()
    psvwave_b200.forward_batch(models, 0, False)
    wall = time.perf_counter() - t0
    dev = lead._get_stat("elapsed_ms") * 1e-3
    n_seq = min(B, 8)
    models[0].forward_simulate(0, False)
    t0 = time.perf_counter()
    seq_dev = 0.0
    for m in models[:n_seq]:
        m.forward_simulate(0, False)
        seq_dev += m._get_stat("elapsed_ms") * 1e-3
    seq_wall = time.perf_counter() - t0
    print(json.dumps({
        "workload": "C4-monte-carlo-forward", "models": B, "grid": [lead.nx, lead.nz], "nt": nt, "options": opts,
        "batched": {"models_per_s": B / dev, "updates_per_s": B * pts * nt / dev, "device_s": dev, "wall_s": wall,
                    "launches": lead._get_stat("launches")},
        "one_by_one": {"models_per_s": n_seq / seq_dev, "updates_per_s": n_seq * pts * nt / seq_dev,
                       "device_s_per_model": seq_dev / n_seq, "wall_s_per_model": seq_wall / n_seq, "sampled": n_seq},
    }), flush=True)


if __name__ == "__main__":
    main()
