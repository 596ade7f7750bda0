"""Variant 4 (two time steps per launch, pair.cuh) on the bench grid: forward (no store) per-step time for
pair_warps x pair_strip, next to the tuned default (run under gpurun, one GPU)."""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench, psvwave_b200
cfg = bench.workload_config(400, 2)
ini = os.path.join(tempfile.mkdtemp(), "w.ini"); bench.write_ini(ini, cfg)
med = bench.marmousi_like(cfg["nx_inner"] + 100, cfg["nz_inner"] + 100)
m = psvwave_b200.fdModel(ini)
m.set_parameter_fields(*med)
pts = m.nx * m.nz
def run(opts):
    for k, v in opts.items(): m._set_option(k, v)
    m.forward_simulate(0, False); m.forward_simulate(0, False)
    ms = m._get_stat("elapsed_ms")
    print(json.dumps({"options": opts, "us_per_step": ms * 1e3 / m.nt, "G_updates_per_s": pts * m.nt / ms / 1e6,
                      "pair_launches": m._get_stat("pair_launches")}), flush=True)
run({"variant": 3})
for w in (6, 4):
    for strip in (0, 32, 48, 64, 96, 128, 192, 256, 400):
        run({"variant": 4, "pair_warps": w, "pair_strip": strip})
