"""One forward march of C1 (default 250x150 grid) launched kernel by kernel: the target of an ncu launch list."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200  # noqa: E402
from helpers import C1_INI  # noqa: E402

m = psvwave_b200.fdModel(C1_INI)
m._set_option("graph", 0)
m.forward_simulate(0, True)
print("elapsed_ms", m._get_stat("elapsed_ms"))
