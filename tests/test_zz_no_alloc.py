"""psv_l2_misfit_static_scratch against psv_calculate_l2_misfit, both through the C ABI (ctypes).

The entry point under test serves programs that are not allowed to allocate GPU memory while they run -- a host
application whose own arena owns the whole card, or a sampling loop that cannot take the device-wide stall of a
cudaMalloc.  Its partial sums therefore sit in a fixed array inside the library instead of a buffer the model grows on
demand.  That array is shared and never cleared, so the test makes two calls in a row, a long one (38 block partials)
followed by a short one (1 block partial): the second result must not see what the first left behind.  Both values are
compared with the handle-based entry point on the same seismograms; the arithmetic and the order of additions are the
same by construction, so the comparison is exact (tolerance 0; the suite's misfit tolerance elsewhere is 1e-14).

Device memory for the inputs comes from torch here -- in the test, not in the library: the entry point itself calls no
allocator.  The file name sorts last on purpose: `pytest -x` stops at the first failure and this test is not allowed
to hide the others.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import C1_INI, SMALL_INI
from helpers import array_ids

pytestmark = pytest.mark.gpu

MAX_SAMPLES = 33554432  # PSV_MISFIT_STATIC_MAX_SAMPLES (include/psvwave_b200.h)
PSV_REAL_DT = 2
PSV_ERR_INVALID_ARGUMENT, PSV_ERR_LENGTH = 1, 2


def library():
    import psvwave_b200
    lib = C.CDLL(psvwave_b200.LIBRARY_PATH)
    lib.psv_last_error.restype = C.c_char_p
    lib.psv_l2_misfit_static_scratch.argtypes = [C.c_void_p] * 4 + [C.c_size_t, C.c_double, C.c_void_p, C.c_void_p]
    lib.psv_l2_misfit_static_scratch.restype = C.c_int
    return lib


def seismograms(lib, ini, seed):
    """A model whose synthetic and observed seismograms are seeded noise: handle, the four host arrays, dt."""
    h = C.c_void_p()
    assert lib.psv_create_from_ini(ini.encode(), C.byref(h)) == 0, lib.psv_last_error()
    ids = array_ids()
    n = C.c_size_t()
    assert lib.psv_array_size(h, ids["rtf_ux"], C.byref(n)) == 0
    rng = np.random.default_rng(seed)
    data = {k: rng.standard_normal(n.value) for k in ("rtf_ux", "rtf_uz", "rtf_ux_true", "rtf_uz_true")}
    for k, a in data.items():
        assert lib.psv_set_array(h, ids[k], a.ctypes.data_as(C.POINTER(C.c_double)), C.c_size_t(a.size)) == 0
    dt = C.c_double()
    assert lib.psv_get_real(h, PSV_REAL_DT, C.byref(dt)) == 0
    return h, data, dt.value


def static_scratch_misfit(lib, torch, data, first, count, dt):
    dev = [torch.from_numpy(data[k][first:first + count]).to("cuda:0")
           for k in ("rtf_ux", "rtf_uz", "rtf_ux_true", "rtf_uz_true")]
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    with torch.cuda.device(0):
        rc = lib.psv_l2_misfit_static_scratch(*(t.data_ptr() for t in dev), count, dt, out.data_ptr(), None)
    assert rc == 0, lib.psv_last_error()
    torch.cuda.synchronize()
    return float(out.cpu()[0])


def test_static_scratch_misfit_matches_allocating_entry_point_twice_in_a_row():
    import torch
    lib = library()
    want = C.c_double()

    # call 1: the reference's default configuration, 1 x 19 x 8000 samples = 38 block partials
    h1, d1, dt1 = seismograms(lib, C1_INI, seed=31)
    assert lib.psv_calculate_l2_misfit(h1, C.byref(want)) == 0, lib.psv_last_error()
    n1 = d1["rtf_ux"].size
    assert n1 > 4096 * 30
    got = static_scratch_misfit(lib, torch, d1, 0, n1, dt1)
    assert want.value > 0 and got == want.value

    # call 2, straight after: one shot of the small two-shot model, 4 x 400 samples = 1 block partial; slots 1..37
    # of the shared array still hold call 1's sums
    h2, d2, dt2 = seismograms(lib, SMALL_INI, seed=32)
    per_shot = d2["rtf_ux"].size // 2
    assert lib.psv_calculate_l2_misfit_shot(h2, 1, C.byref(want)) == 0, lib.psv_last_error()
    got = static_scratch_misfit(lib, torch, d2, per_shot, per_shot, dt2)
    assert want.value > 0 and got == want.value

    # and the long one again: same value as before
    assert lib.psv_calculate_l2_misfit(h1, C.byref(want)) == 0
    assert static_scratch_misfit(lib, torch, d1, 0, n1, dt1) == want.value

    # an empty range is 0.0, like the handle-based call on a model without receivers
    assert static_scratch_misfit(lib, torch, d2, 0, 0, dt2) == 0.0

    assert lib.psv_destroy(h1) == 0 and lib.psv_destroy(h2) == 0


def test_static_scratch_misfit_rejects_what_it_cannot_hold():
    import torch
    lib = library()
    x = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    p = x.data_ptr()
    rc = lib.psv_l2_misfit_static_scratch(p, p, p, p, MAX_SAMPLES + 1, 1.0, p, None)  # rejected before any launch
    assert rc == PSV_ERR_LENGTH
    assert b"PSV_MISFIT_STATIC_MAX_SAMPLES" in lib.psv_last_error()
    assert lib.psv_l2_misfit_static_scratch(p, p, p, p, 8, 1.0, None, None) == PSV_ERR_INVALID_ARGUMENT
    assert lib.psv_l2_misfit_static_scratch(None, p, p, p, 8, 1.0, p, None) == PSV_ERR_INVALID_ARGUMENT
    torch.cuda.synchronize()
    assert float(x.abs().sum().cpu()) == 0.0  # nothing was launched on the rejected calls
