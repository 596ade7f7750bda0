"""ncu target: the resident whole-loop kernel on the default 250x150 grid, nt = 400 (one forward with snapshots)."""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200
from helpers import make_config
ini = os.path.join(tempfile.mkdtemp(), "c1.ini")
make_config(ini, nt=400)
m = psvwave_b200.fdModel(ini)
m._set_option("resident", 1)
m.forward_simulate(0, True)
m.forward_simulate(0, True)
print("us/step", m._get_stat("elapsed_ms") * 1e3 / m.nt)
