"""Single-model forward on the default 250x150 grid (C1) for several launch geometries (run under gpurun)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200
from helpers import C1_INI
for opts in ({}, {"fused_strip": 2}, {"fused_strip": 3}, {"fused_strip": 4}, {"fused_strip": 6}, {"pipe_stages": 2, "fused_strip": 2},
             {"pipe_stages": 2, "fused_strip": 4}, {"variant": 2}, {"variant": 2, "fused_strip": 2}, {"variant": 2, "fused_strip": 4},
             {"variant": 2, "fused_strip": 4, "fused_block": 64}, {"variant": 2, "fused_strip": 2, "fused_block": 32},
             {"variant": 2, "fused_strip": 4, "fused_block": 32}, {"variant": 2, "fused_strip": 8, "fused_block": 32}):
    m = psvwave_b200.fdModel(C1_INI)
    for k, v in opts.items():
        m._set_option(k, v)
    m.forward_simulate(0, False)
    m.forward_simulate(0, False)
    ms = m._get_stat("elapsed_ms")
    print(opts, "us/step %.2f" % (ms * 1e3 / m.nt), "Gupd/s %.2f" % (m.nx * m.nz * m.nt / ms / 1e6), flush=True)
