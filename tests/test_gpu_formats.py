"""What the C ABI exports next to the time loop, checked against the compiled UNMODIFIED reference (oracle/_ref):

  * `psv_create(const psv_config*)` -- the reference's explicit 30-argument constructor (src/fdModel.h:50-64) --
    against `psv_create_from_ini` and against the reference built both ways: the reference's own
    tests/test_constructor_comparison.cpp:14-104 and tests/test_variable_constructor.cpp:22-122 re-expressed;
  * the text formats: `write_receivers` / `write_sources` (src/fdModel.cpp:971-1032), `write_kernels` (:1326-1356)
    and the `output_wavefields` dumps (:753-776) BYTE-compared with the files the reference writes from the same
    state ("exact" arithmetic, so the numbers are bit-identical and the formatting is what is under test);
  * `load_model` (:1185-1289) against the reference reading the same three text files.
"""
import ctypes as C
import filecmp
import os

import numpy as np
import pytest

from helpers import array_ids, gpu_array, gradient_workflow, make_config, ragged_config, random_medium
from oracle import oracle
from oracle.oracle import FullConfig, RefModel, config_from_ini, full_config

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")]

INT_KEYS = ("nt nx nz nx_inner nz_inner nx_inner_boundary nz_inner_boundary np_boundary n_sources n_shots nr "
            "snapshot_interval snapshots nx_free_parameters nz_free_parameters free_parameters basis_gridpoints_x "
            "basis_gridpoints_z").split()
REAL_KEYS = "dx dz dt np_factor scalar_rho scalar_vp scalar_vs peak_frequency alpha t0 delay_cycles_per_shot".split()


class CHandle:
    """A solver object driven through the C ABI only (ctypes), so that psv_create can be reached."""

    def __init__(self, cfg=None, ini=None):
        import psvwave_b200
        lib = C.CDLL(psvwave_b200.LIBRARY_PATH)
        lib.psv_last_error.restype = C.c_char_p
        lib.psv_create.argtypes = [C.POINTER(FullConfig), C.POINTER(C.c_void_p)]
        self.lib, self.h, self.ids = lib, C.c_void_p(), array_ids()
        if cfg is not None:
            c, self._keep = full_config(cfg)
            rc = lib.psv_create(C.byref(c), C.byref(self.h))
        else:
            rc = lib.psv_create_from_ini(ini.encode(), C.byref(self.h))
        if rc != 0:
            raise ValueError((rc, lib.psv_last_error().decode()))

    def __del__(self):
        if self.h:
            self.lib.psv_destroy(self.h)

    def ok(self, rc):
        assert rc == 0, self.lib.psv_last_error()

    def integer(self, key):
        v = C.c_longlong()
        self.ok(self.lib.psv_get_int(self.h, key, C.byref(v)))
        return v.value

    def real(self, key):
        v = C.c_double()
        self.ok(self.lib.psv_get_real(self.h, key, C.byref(v)))
        return v.value

    def array(self, name):
        n = C.c_size_t()
        self.ok(self.lib.psv_array_size(self.h, self.ids[name], C.byref(n)))
        a = np.empty(n.value)
        self.ok(self.lib.psv_get_array(self.h, self.ids[name], a.ctypes.data_as(C.POINTER(C.c_double)), n))
        return a

    def index_list(self, which):
        n = C.c_size_t()
        self.ok(self.lib.psv_index_list_size(self.h, which, C.byref(n)))
        a = np.empty(n.value, dtype=np.int32)
        self.ok(self.lib.psv_get_index_list(self.h, which, a.ctypes.data_as(C.POINTER(C.c_int)), n))
        return a

    def set_medium(self, vp, vs, rho):
        dp = lambda a: np.ascontiguousarray(a).ctypes.data_as(C.POINTER(C.c_double))  # noqa: E731
        self.ok(self.lib.psv_set_parameter_fields(self.h, dp(vp), dp(vs), dp(rho), C.c_size_t(vp.size)))

    def option(self, name, v):
        self.ok(self.lib.psv_set_option(self.h, name.encode(), C.c_longlong(v)))

    def forward(self, shot, store=True):
        self.ok(self.lib.psv_forward_simulate(self.h, shot, int(store), 0, 0))


# ---- constructors ----------------------------------------------------------------------------------------------
def test_explicit_constructor_equals_ini_constructor(tmp_path):
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    cfg = config_from_ini(ini)
    a, b = CHandle(cfg=cfg), CHandle(ini=ini)
    ra, rb = RefModel(cfg), RefModel(ini)  # the reference, both constructors
    # every scalar the reference's comparison test checks (tests/test_constructor_comparison.cpp:21-60)
    for k, name in enumerate(INT_KEYS):
        assert a.integer(k) == b.integer(k) == getattr(ra, name) == getattr(rb, name), name
    for k, name in enumerate(REAL_KEYS):
        assert a.real(k) == b.real(k) == getattr(ra, name) == getattr(rb, name), name
    for which, name in enumerate(("ix_sources", "iz_sources", "ix_receivers", "iz_receivers")):
        assert np.array_equal(a.index_list(which), b.index_list(which)), name
        assert np.array_equal(a.index_list(which), ra.int_array(name)), name  # np_boundary added once
    assert np.array_equal(a.index_list(4), b.index_list(4)) and np.array_equal(a.index_list(5), b.index_list(5))
    # host tables + homogeneous medium (tests/test_constructor_comparison.cpp:62-104 compares the arrays)
    for name in ("taper", "stf", "moment", "t", "vp", "vs", "rho", "lm", "la", "mu", "b_vx", "b_vz"):
        x, y = a.array(name), b.array(name)
        assert np.array_equal(x, y), name
        assert np.array_equal(x, ra.array(name).ravel()) and np.array_equal(x, rb.array(name).ravel()), name
    # and they simulate the same thing (tests/test_variable_constructor.cpp:95-122: deterministic forward runs)
    vp, vs, rho = random_medium(ra.nx, ra.nz, seed=3)
    for m in (a, b):
        m.set_medium(vp, vs, rho)
        m.option("exact", 1)
    ra.set_parameter_fields(vp, vs, rho)
    for s in range(ra.n_shots):
        a.forward(s)
        b.forward(s)
        ra.forward_simulate(s, True)
    for name in ("rtf_ux", "rtf_uz", "accu_txz", "vx", "tzz"):
        assert np.array_equal(a.array(name), b.array(name)), name
        assert np.array_equal(a.array(name), ra.array(name).ravel()), name


def test_explicit_constructor_reports_bad_configurations(tmp_path):
    """Where the reference aborts on an assert (src/fdModel.cpp:404: free window not a multiple of the basis blocks)
    the C ABI returns PSV_ERR_INVALID_ARGUMENT, from both constructors."""
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    cfg = config_from_ini(ini)
    bad = dict(cfg, nx_inner=cfg["nx_inner"] + 1)
    with pytest.raises(ValueError) as e:
        CHandle(cfg=bad)
    assert e.value.args[0][0] == 1
    bad_ini = str(tmp_path / "bad.ini")
    oracle.write_ini(bad_ini, bad)
    with pytest.raises(ValueError) as e:
        CHandle(ini=bad_ini)
    assert e.value.args[0][0] == 1


# ---- text formats, byte for byte ---------------------------------------------------------------------------------
def both(tmp_path, **extra):
    """The CUDA model and the reference on the same configuration, each writing into its own folder."""
    import psvwave_b200
    out = []
    for who in ("gpu", "ref"):
        d = tmp_path / who
        (d / "snapshots").mkdir(parents=True)
        ini = str(d / "case.ini")
        ragged_config(ini, nt=60, observed_data_folder=str(d), stf_folder=str(d), **extra)
        out.append((psvwave_b200.fdModel(ini) if who == "gpu" else RefModel(ini), d))
    (m, dm), (r, dr) = out
    m._set_option("exact", 1)
    medium = random_medium(r.nx, r.nz, seed=21)
    m.set_parameter_fields(*medium)
    r.set_parameter_fields(*medium)
    return m, dm, r, dr


def same_file(a, b):
    assert os.path.getsize(a) > 0, a
    assert filecmp.cmp(a, b, shallow=False), f"{a} differs from {b}"


def test_receiver_source_and_kernel_files_match_the_reference_byte_for_byte(tmp_path, monkeypatch):
    m, dm, r, dr = both(tmp_path)
    for x in (m, r):
        for s in range(x.n_shots):
            x.forward_simulate(s, True)
    zero = np.zeros((r.n_shots, r.nr, r.nt))
    obs = [a * f for a, f in zip(r.get_synthetic_data(), (1.1, 0.9))]
    for x, d in ((m, dm), (r, dr)):
        x.set_observed_data(*obs)
        gradient_workflow(x)
        monkeypatch.chdir(d)  # write_kernels writes into the current directory
        x.write_receivers()
        x.write_sources()
        x.write_kernels()
    del zero
    for s in range(r.n_shots):
        same_file(dm / f"rtf_ux{s}.txt", dr / f"rtf_ux{s}.txt")
        same_file(dm / f"rtf_uz{s}.txt", dr / f"rtf_uz{s}.txt")
        same_file(dm / f"sources_shot_{s}.txt", dr / f"sources_shot_{s}.txt")
    for k in ("kernel_vp.txt", "kernel_vs.txt", "kernel_density.txt"):
        same_file(dm / k, dr / k)
    # and each side reads the other's receiver files
    monkeypatch.chdir(dm)
    for f in os.listdir(dr):
        if f.startswith("rtf_"):
            os.replace(dr / f, dm / f)
    m.load_receivers()
    for a, b in zip(m.get_observed_data(), r.get_synthetic_data()):
        assert np.array_equal(a, b)


def test_wavefield_dumps_match_the_reference_byte_for_byte(tmp_path, monkeypatch):
    m, dm, r, dr = both(tmp_path)
    for x, d in ((m, dm), (r, dr)):
        monkeypatch.chdir(d)  # the dumps go to ./snapshots/ (src/fdModel.cpp:756-757)
        x.forward_simulate(0, True, False, True)
    names = sorted(os.listdir(dr / "snapshots"))
    assert names == sorted(os.listdir(dm / "snapshots")) and len(names) == 2 * 6  # it = 0, 10, .., 50: vx + vz
    for n in names:
        same_file(dm / "snapshots" / n, dr / "snapshots" / n)
    for a, b in zip(m.get_synthetic_data(), r.get_synthetic_data()):
        assert np.array_equal(a[0], b[0])


def test_load_model_matches_the_reference(tmp_path, monkeypatch, capfd):
    m, dm, r, dr = both(tmp_path)
    vp, vs, rho = random_medium(r.nx, r.nz, seed=33)
    monkeypatch.chdir(tmp_path)
    for name, a in (("de.txt", rho), ("vp.txt", vp), ("vs.txt", vs)):
        np.savetxt(name, a, fmt="%.17g")  # ix-major, iz fastest: one row per ix
    m.load_model("de.txt", "vp.txt", "vs.txt", False)
    r.load_model("de.txt", "vp.txt", "vs.txt", False)
    capfd.readouterr()  # the reference prints one line per grid point (src/fdModel.cpp:1241)
    for name in ("vp", "vs", "rho", "lm", "la", "mu", "b_vx", "b_vz"):
        assert np.array_equal(gpu_array(m, name), r.array(name)), name
    assert np.array_equal(gpu_array(m, "vp"), vp)
    np.savetxt("short.txt", vp[:-1], fmt="%.17g")
    with pytest.raises(ValueError):
        m.load_model("de.txt", "short.txt", "vs.txt", False)
    with pytest.raises(ValueError):
        r.load_model("de.txt", "short.txt", "vs.txt", False)
    with pytest.raises(ValueError):
        m.load_model("de.txt", "missing.txt", "vs.txt", False)
