"""One short pass over everything the default step kernel does, for compute-sanitizer (scripts/sanitize.sh):
plain / snapshot / correlation launches of variant 3 (TMA ring + mbarriers), a batch of 3 models (blockIdx.z),
a 2-slab split with halo exchange, and -- when built -- the resident whole-loop kernel.  Results are checked against
the CPU oracle so that a sanitizer-clean run is also a correct one."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psvwave_b200  # noqa: E402
from oracle.oracle import OracleModel, config_from_ini, write_ini  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
nt = int(os.environ.get("PSV_SANITIZE_NT", "24"))
gold = np.load(os.path.join(ROOT, "tests", "golden", "small_multishot.npz"))
cfg = dict(config_from_ini(os.path.join(ROOT, "tests", "golden", "small_multishot.ini")), nt=nt, snapshot_interval=4)
ini = os.path.join(tempfile.mkdtemp(prefix="psv_san_"), "case.ini")
write_ini(ini, cfg)
medium = (gold["vp"], gold["vs"], gold["rho"])


def workflow(m, shots=(0, 1)):
    m.set_parameter_fields(*medium)
    for s in shots:
        m.forward_simulate(s, True)
    zero = np.zeros((m.n_shots, m.nr, m.nt))
    m.set_observed_data(zero, zero)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    for s in shots:
        m.adjoint_simulate(s)
    m.map_kernels_to_velocity()
    return m.get_synthetic_data()[0], m.get_gradient_vector()


o_ux, o_g = workflow(OracleModel(ini))


def check(tag, ux, g):
    assert np.array_equal(ux, o_ux) and np.array_equal(g, o_g), tag
    print("ok:", tag, flush=True)


def fd(**opts):
    m = psvwave_b200.fdModel(ini)
    m._set_option("exact", 1)
    m._set_option("graph", 0)
    for k, v in opts.items():
        m._set_option(k, v)
    return m


def have_resident():
    try:
        fd(resident=0)
        return True
    except ValueError:
        return False


RESIDENT = have_resident()
off = dict(resident=0) if RESIDENT else {}


if what in ("all", "pipe"):
    check("variant 3 (TMA ring): plain + snapshot + correlation", *workflow(fd(variant=3, **off)))
    check("variant 3, 8 warps, 3 stages", *workflow(fd(variant=3, pipe_warps=8, pipe_stages=3, **off)))
if what in ("all", "batch"):
    ms = [fd(variant=3, **off) for _ in range(3)]
    for m in ms:
        m.set_parameter_fields(*medium)
    for s in (0, 1):
        psvwave_b200.forward_batch(ms, s, True)
    for m in ms:
        zero = np.zeros((m.n_shots, m.nr, m.nt))
        m.set_observed_data(zero, zero)
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
    for s in (0, 1):
        psvwave_b200.adjoint_batch(ms, s)
    for m in ms:
        m.map_kernels_to_velocity()
        check("batch of 3 models", m.get_synthetic_data()[0], m.get_gradient_vector())
if what in ("all", "slab"):
    d = psvwave_b200.SlabDomain(ini, n_slabs=2, halo_steps=1, options=dict(exact=1, variant=3, **off))
    ux, g = workflow(d)
    check("2 slabs, exchange every step", ux, g)
if what in ("all", "resident") and RESIDENT:
    check("resident whole-loop kernel", *workflow(fd(resident=1)))
if what in ("all", "pair"):
    check("variant 4 (two steps per launch)", *workflow(fd(variant=4, **off)))
print("sanitize_case done:", what)
