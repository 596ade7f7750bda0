"""CPU-only checks of the boundary and the host logic: the C-ABI library loads and exports
every symbol include/psvwave_b200.h declares, the .ini reader behaves like the reference's
(ext/inih/INIReader.h + src/fdModel.cpp:470-523,1574-1659), errors map to the reference's
exception types, and there is no CPU fallback."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, SMALL_INI
from helpers import C1_INI, make_config, ragged_config
from oracle.oracle import RefModel, config_from_ini, ref_available

HEADER = os.path.join(ROOT, "include", "psvwave_b200.h")


@pytest.fixture(scope="module")
def lib():
    import psvwave_b200
    return C.CDLL(psvwave_b200.LIBRARY_PATH)


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(psv_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(lib):
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/psvwave_b200.h but not exported"


def check_ini(lib, path):
    buf = C.create_string_buffer(1 << 20)
    lib.psv_check_ini.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    lib.psv_last_error.restype = C.c_char_p
    rc = lib.psv_check_ini(path.encode(), buf, len(buf))
    if rc != 0:
        return rc, lib.psv_last_error().decode()
    out = {}
    for line in buf.value.decode().splitlines():
        k, _, v = line.partition("=")
        out[k] = v
    return 0, out


def ints(s):
    return [int(x) for x in s.split(",")] if s else []


@pytest.mark.parametrize("path", [C1_INI, SMALL_INI])
def test_ini_reader_matches_restated_reference_reader(lib, path, capfd):
    rc, got = check_ini(lib, path)
    assert rc == 0, got
    out = capfd.readouterr().out  # the reference's two progress lines (src/fdModel.cpp:396,472)
    assert f"Loading configuration file: '{path}'." in out and "Parsing passed configuration." in out
    want = config_from_ini(path)
    for k in ("nt", "nx_inner", "nz_inner", "nx_inner_boundary", "nz_inner_boundary", "np_boundary", "npx",
              "npz", "n_sources", "n_shots", "nr", "snapshot_interval"):
        assert int(got[k]) == want[k], k
    for k in ("dx", "dz", "dt", "np_factor", "scalar_rho", "scalar_vp", "scalar_vs", "peak_frequency",
              "source_timeshift", "delay_cycles_per_shot"):
        assert float(got[k]) == want[k], k
    npb = want["np_boundary"]
    assert ints(got["ix_sources"]) == [v + npb for v in want["ix_sources"]]
    assert ints(got["iz_receivers"]) == [v + npb for v in want["iz_receivers"]]
    assert [float(x) for x in got["moment_angles"].split(",")] == want["moment_angles"]
    assert [ints(s) for s in got["which_source_to_fire_in_which_shot"].split("|")] == \
        want["which_source_to_fire_in_which_shot"]
    assert int(got["nx"]) == want["nx_inner"] + 2 * npb
    assert int(got["free_parameters"]) == 3 * (want["nx_inner"] - 2 * want["nx_inner_boundary"]) * \
        (want["nz_inner"] - 2 * want["nz_inner_boundary"]) // (want["npx"] * want["npz"])
    assert int(got["snapshots"]) == want["nt"] // want["snapshot_interval"]


@pytest.mark.skipif(not ref_available(), reason="reference core not built here")
def test_ini_reader_matches_live_reference(lib, tmp_path):
    path = str(tmp_path / "r.ini")
    ragged_config(path)
    rc, got = check_ini(lib, path)
    assert rc == 0, got
    r = RefModel(path)
    for k in ("nt", "nx", "nz", "np_boundary", "n_sources", "n_shots", "nr", "snapshots", "free_parameters",
              "nx_free_parameters", "nz_free_parameters"):
        assert int(got[k]) == getattr(r, k), k
    for k in ("ix_sources", "iz_sources", "ix_receivers", "iz_receivers"):
        assert ints(got[k]) == list(r.int_array(k)), k


def test_ini_quirks(lib, tmp_path):
    text = open(C1_INI).read()
    # (1) a 19-receiver list pushed far beyond the reference's 200-character line limit,
    #     (2) a continuation line, (3) ':' separator, (4) upper-case section/key, (5) hex integer
    many = ", ".join(str(10 + (i % 180)) for i in range(400))
    zs = ", ".join("90" for _ in range(400))
    text = re.sub(r"nr = 19;.*", "nr = 400", text)
    text = re.sub(r"ix_receivers = .*", "ix_receivers = {" + many + "}", text)
    text = re.sub(r"iz_receivers = .*", "iz_receivers = {" + zs[: len(zs) // 2] + "\n   " + zs[len(zs) // 2:] + "}", text)
    text = text.replace("[boundary]", "[BOUNDARY]").replace("np_factor = 0.015;", "NP_Factor : 0.015 ; comment")
    text = text.replace("nt = 8000;", "nt = 0x1F40;")
    p = tmp_path / "quirks.ini"
    p.write_text(text)
    rc, got = check_ini(lib, str(p))
    assert rc == 0, got
    assert int(got["nr"]) == 400 and len(ints(got["ix_receivers"])) == 400 and len(ints(got["iz_receivers"])) == 400
    assert float(got["np_factor"]) == 0.015 and int(got["nt"]) == 8000


@pytest.mark.parametrize("mutate,code,needle", [
    (lambda t: t.replace("dt = 0.00025;", ""), 1, "(domain=dt) does not exist"),
    (lambda t: t.replace("nt = 8000;", "nt = many;"), 1, "does not seem to be an integer"),
    (lambda t: t.replace("n_sources = 4;", "n_sources = 3;"), 1, "Dimension mismatch between n_sources"),
    (lambda t: t.replace("n_shots = 1;", "n_shots = 2;"), 1, "Mismatch between n_shots"),
    (lambda t: t.replace("{{0, 1, 2, 3}}", "{{0, 1, 2}}"), 1, "Mismatch between n_sources and sources.which"),
    (lambda t: t.replace("nr = 19;", "nr = 18;"), 1, "Mismatch between nr and receivers"),
])
def test_ini_errors_match_reference_messages(lib, tmp_path, mutate, code, needle):
    p = tmp_path / "bad.ini"
    p.write_text(mutate(open(C1_INI).read()))
    rc, msg = check_ini(lib, str(p))
    assert rc == code and needle in msg, (rc, msg)
    if ref_available():  # the reference throws std::invalid_argument with the same text
        with pytest.raises(ValueError) as e:
            RefModel(str(p))
        assert needle in str(e.value)


def test_missing_file(lib, tmp_path):
    rc, msg = check_ini(lib, str(tmp_path / "nope.ini"))
    assert rc == 1 and "Can't load .ini file." in msg  # src/fdModel.cpp:476-479


def test_python_surface_matches_reference_binding():
    """Every method / attribute the reference binds (src/psvWave.cpp:465-655) exists."""
    import psvwave_b200 as p
    methods = ("copy forward_simulate adjoint_simulate get_model_vector set_model_vector get_gradient_vector "
               "load_vector get_snapshots get_extent get_coordinates get_parameter_fields set_parameter_fields "
               "get_kernels get_synthetic_data get_observed_data set_synthetic_data set_observed_data "
               "get_sources get_receivers calculate_l2_misfit calculate_l2_adjoint_sources reset_kernels "
               "map_kernels_to_velocity").split()
    attrs = ("n_sources dt dz dx nt nz nx nx_free_parameters nz_free_parameters nx_inner nz_inner "
             "nx_inner_boundary nz_inner_boundary np_boundary free_parameters snapshot_interval snapshots "
             "n_shots which_source_to_fire_in_which_shot misfit").split()
    for name in methods + attrs:
        assert hasattr(p.fdModel, name), name
    assert p.fdModel.units == [r"$m/s$", r"$m/s$", r"$kg/m^3$"]
    d = p.get_dictionary()
    assert set(d) == {"domain", "boundary", "medium", "sources", "receivers", "inversion", "basis", "output"}


def test_write_dictionary_round_trips_through_the_native_reader(lib, tmp_path):
    import psvwave_b200 as p
    d = p.get_dictionary()
    d["sources"]["which_source_to_fire_in_which_shot"] = [[0], [1]]
    d["sources"]["n_shots"] = 2
    path = str(tmp_path / "auto.conf")
    p.write_dictionary(path, d)
    rc, got = check_ini(lib, path)
    assert rc == 0, got
    assert int(got["nx"]) == 450 and got["which_source_to_fire_in_which_shot"] == "0|1"
    assert ints(got["ix_sources"]) == [175, 375]


def test_no_cpu_fallback():
    """Without a CUDA device construction must fail loudly (never compute on the host)."""
    import psvwave_b200 as p
    if p.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        p.fdModel(C1_INI)


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under psvwave_b200/ may name it."""
    pkg = os.path.join(ROOT, "psvwave_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower() or f == "parallel.py" and False, os.path.join(dirpath, f)


@pytest.mark.parametrize("sigma", [0.5, 0.7, 1.0, 2.0, 2.5, 4.0, 7.3])
def test_gaussian_weights_are_scipys(lib, sigma):
    """The filter weights of psv_get_gradient_vector_smoothed (host code, no device): scipy.ndimage's
    _gaussian_kernel1d(sigma, 0, int(4 sigma + 0.5)), the kernel behind gaussian_filter(..., sigma) that the
    reference's _fwi class applies to the gradient (docs-source/source/typical-workflow.rst:61-69)."""
    import numpy as np
    from scipy.ndimage import _filters
    radius = C.c_int(0)
    buf = (C.c_double * 4096)()
    lib.psv_gaussian_weights.argtypes = [C.c_double, C.POINTER(C.c_double), C.c_size_t,
                                         C.POINTER(C.c_int)]
    assert lib.psv_gaussian_weights(sigma, buf, 4096, C.byref(radius)) == 0
    lw = int(4.0 * sigma + 0.5)
    assert radius.value == lw
    want = _filters._gaussian_kernel1d(sigma, 0, lw)
    got = np.array(buf[:2 * lw + 1])
    assert np.allclose(got, want, rtol=4e-16, atol=0)  # libm exp vs numpy's SIMD exp: at most an ulp or two
    assert abs(got.sum() - 1.0) <= 4e-16
    assert lib.psv_gaussian_weights(-1.0, buf, 4096, C.byref(radius)) != 0
    assert lib.psv_gaussian_weights(sigma, buf, 1, C.byref(radius)) != 0  # capacity too small


def test_reference_import_names_resolve_to_this_library():
    """`import psvWave` / top-level `__psvWave_cpp` (psvWave/__init__.py:6 of the reference) are this library."""
    import __psvWave_cpp
    import psvWave
    import psvwave_b200
    assert psvWave.fdModel is psvwave_b200.fdModel is __psvWave_cpp.fdModel
    assert psvWave.get_dictionary()["domain"]["nt"] == psvwave_b200.get_dictionary()["domain"]["nt"]
    assert os.path.exists(os.path.join(ROOT, "tests", "test_configurations", "default_testing_configuration.ini"))
