"""Debug driver: small multishot forward with variant 5 vs variant 3 (bit-exact arithmetic)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psvwave_b200
ini = os.path.join(ROOT, "tests", "golden", "small_multishot.ini")
nt_opt = {}
def run(variant, **opts):
    m = psvwave_b200.fdModel(ini)
    m._set_option("variant", variant); m._set_option("exact", 1)
    for k, v in opts.items(): m._set_option(k, v)
    m.forward_simulate(0, True)
    ux, uz = m.get_synthetic_data()
    snaps = m.get_snapshots()
    return ux[0].copy(), uz[0].copy(), [s[0].copy() for s in snaps], m
a = run(3)
print("v3 done", flush=True)
b = run(5)
print("v5 done launches", b[3]._get_stat("launches"), "pairs", b[3]._get_stat("pair_launches"), flush=True)
print("ux equal", np.array_equal(a[0], b[0]), "uz equal", np.array_equal(a[1], b[1]))
for i, f in enumerate("vx vz txx tzz txz".split()):
    d = a[2][i] - b[2][i]
    bad = np.argwhere(d != 0)
    print(f, "snap equal", bad.size == 0, "first bad (snap, ix, iz)", bad[:3].tolist() if bad.size else None, "n bad", len(bad))
