"""Where a time step of the resident loop kernel goes: clock cycles of thread 0 per phase, mean and max over the CTAs
(option "resident_timing"), on the default 250x150 configuration.  Run under gpurun."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200  # noqa: E402
from helpers import C1_INI  # noqa: E402

PHASES = ["halo: poll + load the neighbours' velocity packets, barrier", "stress rows + barrier",
          "velocity rows + injections + packets out + barrier", "snapshot / correlation / receivers of the next step",
          "-", "-", "-", "-"]
for opts in [dict(a.split("=") for a in arg.split(",") if a) for arg in (sys.argv[1:] or [""])]:
    m = psvwave_b200.fdModel(C1_INI)
    m._set_option("resident", 1)
    m._set_option("resident_timing", 1)
    for k, v in opts.items():
        m._set_option(k, int(v))
    out = {"options": opts}
    for name, call in (("forward", lambda: m.forward_simulate(0, True)), ("adjoint", lambda: m.adjoint_simulate(0))):
        call()
        call()
        out[name] = {"us_per_step": m._get_stat("elapsed_ms") * 1e3 / m.nt, "ctas": m._get_stat("resident_ctas"),
                     "cycles_per_step_mean": {p: round(m._get_stat(f"resident_phase{k}")) for k, p in enumerate(PHASES)},
                     "cycles_per_step_max": {p: round(m._get_stat(f"resident_phase_max{k}")) for k, p in enumerate(PHASES)}}
    print(json.dumps(out), flush=True)
