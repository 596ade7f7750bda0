"""TEST INFRASTRUCTURE ONLY -- ctypes front-ends for the two CPU checkers.

* ``OracleModel``  -> oracle/libpsv_oracle.so, our plain-C restatement (psv_oracle.c).
* ``RefModel``     -> oracle/_ref/libpsv_ref.so, the UNMODIFIED reference core behind
                      oracle/ref_shim.cpp (only where it has been built; it is built in the
                      dev container, where /root/reference exists, and travels to the GPU
                      box as a prebuilt file).

Nothing under psvwave_b200/ imports this module.  Only tests/, __graft_entry__.smoke() and
bench.py (cpu_baseline leg and --impl reference) do.

Both classes expose the same small surface (modelled on the reference's Python class,
src/psvWave.cpp:465-655): forward_simulate, adjoint_simulate, calculate_l2_misfit,
calculate_l2_adjoint_sources, reset_kernels, map_kernels_to_velocity, get/set model vector,
get_gradient_vector, set_parameter_fields, set_observed_data, ... plus ``array(name)`` which
returns a *view* of any internal array under the reference's own field names
(src/fdModel.h:209-252).
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORC_LIB = os.path.join(HERE, "libpsv_oracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libpsv_ref.so")


def build(quiet: bool = True) -> None:
    """Compile the C restatement (always) and oracle/_ref (when /root/reference exists)."""
    subprocess.run(["make", "-C", HERE], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


# --------------------------------------------------------------------------------------
# .ini reading -- a restatement of how the reference reads its configuration
# (ext/inih/INIReader.h:186-283 line rules; src/fdModel.cpp:470-523 keys;
#  src/fdModel.cpp:1574-1659 list syntax).  Used to feed OracleModel; the product has its
#  own C++ parser which tests compare against this one and against RefModel.
# --------------------------------------------------------------------------------------
def _strip_inline_comment(s: str) -> str:
    # a ';' starts a comment only when preceded by whitespace (INIReader.h:140-154)
    prev_space = False
    for i, ch in enumerate(s):
        if prev_space and ch == ";":
            return s[:i]
        prev_space = ch.isspace()
    return s


def read_ini(path: str) -> dict:
    values: dict[str, str] = {}
    section = ""
    prev = ""
    with open(path, "r") as fh:
        for raw in fh:
            line = raw.rstrip()
            start = line.lstrip()
            if start[:1] in (";", "#"):
                continue
            if prev and start and len(start) < len(line):  # continuation line
                v = _strip_inline_comment(start).rstrip()
                key = (section + "=" + prev).lower()
                values[key] = (values[key] + "\n" if values.get(key) else "") + v
            elif start[:1] == "[":
                m = re.match(r"\[([^\]]*)\]", start)
                if not m:
                    break
                section = m.group(1)
                prev = ""
            elif start:
                m = re.search(r"[=:]", start)
                if not m:
                    break
                name = start[: m.start()].rstrip()
                v = _strip_inline_comment(start[m.end():].lstrip()).rstrip()
                prev = name
                key = (section + "=" + name).lower()
                values[key] = (values[key] + "\n" if values.get(key) else "") + v
    return values


def _num_prefix(s: str, integer: bool):
    pat = r"\s*[-+]?(0[xX][0-9a-fA-F]+|\d+)" if integer else \
        r"\s*[-+]?(\d+\.?\d*([eE][-+]?\d+)?|\.\d+([eE][-+]?\d+)?)"
    m = re.match(pat, s)
    if not m:
        raise ValueError("not a number: %r" % s)
    return int(m.group(0), 0) if integer else float(m.group(0))


def _flat_list(s: str):
    s = re.sub(r"\s", "", s)
    s = s[: s.index("}")][1:]
    return [_atof(tok) for tok in s.split(",")]


def _atof(tok: str) -> float:
    m = re.match(r"[-+]?(\d+\.?\d*([eE][-+]?\d+)?|\.\d+([eE][-+]?\d+)?)", tok)
    return float(m.group(0)) if m else 0.0


def _nested_list(s: str):
    s = re.sub(r"\s", "", s)[2:]
    parts = s.split("},{")
    last = parts[-1]
    k = last.find("}};")
    if k >= 0:
        last = last[:k]
    parts[-1] = last
    return [[int(_atof(tok)) for tok in p.split(",")] for p in parts]


def config_from_ini(path: str) -> dict:
    v = read_ini(path)

    def get(sec, name):
        key = (sec + "=" + name).lower()
        if key not in v:
            raise ValueError("The requested key (%s) does not exist in the .ini file (%s)." % (key, path))
        return v[key]

    cfg = dict(
        nt=_num_prefix(get("domain", "nt"), True),
        nx_inner=_num_prefix(get("domain", "nx_inner"), True),
        nz_inner=_num_prefix(get("domain", "nz_inner"), True),
        nx_inner_boundary=_num_prefix(get("domain", "nx_inner_boundary"), True),
        nz_inner_boundary=_num_prefix(get("domain", "nz_inner_boundary"), True),
        dx=_num_prefix(get("domain", "dx"), False),
        dz=_num_prefix(get("domain", "dz"), False),
        dt=_num_prefix(get("domain", "dt"), False),
        np_boundary=_num_prefix(get("boundary", "np_boundary"), True),
        np_factor=_num_prefix(get("boundary", "np_factor"), False),
        scalar_rho=_num_prefix(get("medium", "scalar_rho"), False),
        scalar_vp=_num_prefix(get("medium", "scalar_vp"), False),
        scalar_vs=_num_prefix(get("medium", "scalar_vs"), False),
        npx=_num_prefix(get("basis", "npx"), True),
        npz=_num_prefix(get("basis", "npz"), True),
        peak_frequency=_num_prefix(get("sources", "peak_frequency"), False),
        source_timeshift=_num_prefix(get("sources", "source_timeshift"), False),
        delay_cycles_per_shot=_num_prefix(get("sources", "delay_cycles_per_shot"), False),
        n_sources=_num_prefix(get("sources", "n_sources"), True),
        n_shots=_num_prefix(get("sources", "n_shots"), True),
        ix_sources=[int(x) for x in _flat_list(get("sources", "ix_sources"))],
        iz_sources=[int(x) for x in _flat_list(get("sources", "iz_sources"))],
        moment_angles=_flat_list(get("sources", "moment_angles")),
        which_source_to_fire_in_which_shot=_nested_list(
            get("sources", "which_source_to_fire_in_which_shot")),
        nr=_num_prefix(get("receivers", "nr"), True),
        ix_receivers=[int(x) for x in _flat_list(get("receivers", "ix_receivers"))],
        iz_receivers=[int(x) for x in _flat_list(get("receivers", "iz_receivers"))],
        snapshot_interval=_num_prefix(get("inversion", "snapshot_interval"), True),
        observed_data_folder=get("output", "observed_data_folder"),
        stf_folder=get("output", "stf_folder"),
    )
    return cfg


def write_ini(path: str, cfg: dict) -> None:
    """Write a config dict (as returned by config_from_ini) in the reference's .ini format."""
    def lst(x):
        return "{" + ", ".join(repr(v) if isinstance(v, float) else str(v) for v in x) + "}"

    nested = "{" + ", ".join(lst(s) for s in cfg["which_source_to_fire_in_which_shot"]) + "}"
    lines = [
        "[domain]",
        *(f"{k} = {cfg[k]!r}" for k in ("nt", "nx_inner", "nz_inner", "nx_inner_boundary",
                                        "nz_inner_boundary", "dx", "dz", "dt")),
        "[boundary]", f"np_boundary = {cfg['np_boundary']}", f"np_factor = {cfg['np_factor']!r}",
        "[medium]",
        *(f"{k} = {cfg[k]!r}" for k in ("scalar_rho", "scalar_vp", "scalar_vs")),
        "[sources]",
        *(f"{k} = {cfg[k]!r}" for k in ("peak_frequency", "n_sources", "n_shots",
                                        "source_timeshift", "delay_cycles_per_shot")),
        f"moment_angles = {lst(cfg['moment_angles'])}",
        f"ix_sources = {lst(cfg['ix_sources'])}",
        f"iz_sources = {lst(cfg['iz_sources'])}",
        f"which_source_to_fire_in_which_shot = {nested}",
        "[receivers]", f"nr = {cfg['nr']}",
        f"ix_receivers = {lst(cfg['ix_receivers'])}",
        f"iz_receivers = {lst(cfg['iz_receivers'])}",
        "[inversion]", f"snapshot_interval = {cfg['snapshot_interval']}",
        "[basis]", f"npx = {cfg['npx']}", f"npz = {cfg['npz']}",
        "[output]", f"observed_data_folder = {cfg.get('observed_data_folder', '.')}",
        f"stf_folder = {cfg.get('stf_folder', '.')}",
    ]
    with open(path, "w") as fh:
        fh.write("\n".join(lines) + "\n")


# --------------------------------------------------------------------------------------
class _OrcConfig(C.Structure):
    _fields_ = [
        ("nt", C.c_int), ("nx_inner", C.c_int), ("nz_inner", C.c_int),
        ("nx_inner_boundary", C.c_int), ("nz_inner_boundary", C.c_int),
        ("dx", C.c_double), ("dz", C.c_double), ("dt", C.c_double),
        ("np_boundary", C.c_int), ("np_factor", C.c_double),
        ("scalar_rho", C.c_double), ("scalar_vp", C.c_double), ("scalar_vs", C.c_double),
        ("npx", C.c_int), ("npz", C.c_int),
        ("peak_frequency", C.c_double), ("source_timeshift", C.c_double),
        ("delay_cycles_per_shot", C.c_double),
        ("n_sources", C.c_int), ("n_shots", C.c_int),
        ("ix_sources", C.POINTER(C.c_int)), ("iz_sources", C.POINTER(C.c_int)),
        ("moment_angles", C.POINTER(C.c_double)),
        ("shot_offsets", C.POINTER(C.c_int)), ("shot_sources", C.POINTER(C.c_int)),
        ("nr", C.c_int),
        ("ix_receivers", C.POINTER(C.c_int)), ("iz_receivers", C.POINTER(C.c_int)),
        ("snapshot_interval", C.c_int),
    ]


class FullConfig(C.Structure):
    """include/psvwave_b200.h `psv_config` = oracle/ref_shim.cpp `ref_config`: the reference's explicit
    30-argument constructor (src/fdModel.h:50-64) as one flat record."""
    _fields_ = _OrcConfig._fields_ + [("observed_data_folder", C.c_char_p), ("stf_folder", C.c_char_p)]


def full_config(cfg: dict):
    """(FullConfig, keep-alive list) from a config dict as returned by config_from_ini."""
    which = cfg["which_source_to_fire_in_which_shot"]
    offs = np.zeros(cfg["n_shots"] + 1, dtype=np.int32)
    offs[1:] = np.cumsum([len(s) for s in which])
    keep = [np.array(cfg["ix_sources"], dtype=np.int32), np.array(cfg["iz_sources"], dtype=np.int32),
            np.array(cfg["moment_angles"], dtype=np.float64), offs,
            np.array([s for sub in which for s in sub], dtype=np.int32),
            np.array(cfg["ix_receivers"], dtype=np.int32), np.array(cfg["iz_receivers"], dtype=np.int32)]
    ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))  # noqa: E731
    c = FullConfig()
    for k in ("nt", "nx_inner", "nz_inner", "nx_inner_boundary", "nz_inner_boundary", "dx", "dz", "dt", "np_boundary",
              "np_factor", "scalar_rho", "scalar_vp", "scalar_vs", "npx", "npz", "peak_frequency", "source_timeshift",
              "delay_cycles_per_shot", "n_sources", "n_shots", "nr", "snapshot_interval"):
        setattr(c, k, cfg[k])
    c.ix_sources, c.iz_sources = ip(keep[0]), ip(keep[1])
    c.moment_angles = keep[2].ctypes.data_as(C.POINTER(C.c_double))
    c.shot_offsets, c.shot_sources = ip(keep[3]), ip(keep[4])
    c.ix_receivers, c.iz_receivers = ip(keep[5]), ip(keep[6])
    c.observed_data_folder = str(cfg.get("observed_data_folder", ".")).encode()
    c.stf_folder = str(cfg.get("stf_folder", ".")).encode()
    return c, keep


def _load(path: str, prefix: str):
    lib = C.CDLL(path)
    f = lambda n: getattr(lib, prefix + n)  # noqa: E731
    f("get_int").restype = C.c_long
    f("get_int").argtypes = [C.c_void_p, C.c_char_p]
    f("get_real").restype = C.c_double
    f("get_real").argtypes = [C.c_void_p, C.c_char_p]
    f("array").restype = C.POINTER(C.c_double)
    f("array").argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_longlong)]
    f("int_array").restype = C.POINTER(C.c_int)
    f("int_array").argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_longlong)]
    f("destroy").restype = None
    f("destroy").argtypes = [C.c_void_p]
    for n in ("get_model_vector", "set_model_vector", "get_gradient_vector"):
        f(n).restype = C.c_int
        f(n).argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_longlong]
    return lib


class _CpuModel:
    """Shared behaviour of OracleModel / RefModel."""
    _prefix = ""
    _lib = None
    _h = None

    def _f(self, name):
        return getattr(self._lib, self._prefix + name)

    def __del__(self):
        if self._h and self._lib is not None:
            self._f("destroy")(self._h)
            self._h = None

    def _int(self, name):
        return int(self._f("get_int")(self._h, name.encode()))

    def _real(self, name):
        return float(self._f("get_real")(self._h, name.encode()))

    def __getattr__(self, name):
        # read-only scalars under the reference's attribute names (src/psvWave.cpp:499-626)
        if name.startswith("_"):
            raise AttributeError(name)
        v = self._int(name)
        if v != -999999:
            return v
        r = self._real(name)
        if r != -999999.0:
            return r
        raise AttributeError(name)

    _SHAPES = {
        "grid": ("vx vz txx tzz txz lm la mu b_vx b_vz rho vp vs lambda_kernel mu_kernel "
                 "density_l_kernel vp_kernel vs_kernel density_v_kernel taper").split(),
        "rec": "rtf_ux rtf_uz rtf_ux_true rtf_uz_true a_stf_ux a_stf_uz".split(),
        "accu": "accu_vx accu_vz accu_txx accu_tzz accu_txz".split(),
    }

    def array(self, name: str) -> np.ndarray:
        """A writable VIEW of an internal array (reference field names)."""
        n = C.c_longlong(0)
        p = self._f("array")(self._h, name.encode(), C.byref(n))
        if not p:
            raise KeyError(name)
        a = np.ctypeslib.as_array(p, shape=(n.value,))
        if name in self._SHAPES["grid"]:
            return a.reshape(self.nx, self.nz)
        if name in self._SHAPES["rec"]:
            return a.reshape(self.n_shots, self.nr, self.nt)
        if name in self._SHAPES["accu"]:
            return a.reshape(self.n_shots, self.snapshots, self.nx, self.nz)
        if name == "stf":
            return a.reshape(self.n_sources, self.nt)
        if name == "moment":
            return a.reshape(self.n_sources, 2, 2)
        return a

    def int_array(self, name: str) -> np.ndarray:
        n = C.c_longlong(0)
        p = self._f("int_array")(self._h, name.encode(), C.byref(n))
        if not p:
            raise KeyError(name)
        return np.ctypeslib.as_array(p, shape=(n.value,))

    # -- reference-shaped API ----------------------------------------------------------
    def get_parameter_fields(self):
        return self.array("vp").copy(), self.array("vs").copy(), self.array("rho").copy()

    def set_parameter_fields(self, vp, vs, rho):
        for name, src in (("vp", vp), ("vs", vs), ("rho", rho)):
            dst = self.array(name)
            src = np.ascontiguousarray(src, dtype=np.float64)
            if src.shape != dst.shape:
                raise ValueError("The input ndarray _%s does not have the right shape." % name)
            dst[...] = src
        self.update_from_velocity()

    def get_synthetic_data(self):
        return self.array("rtf_ux").copy(), self.array("rtf_uz").copy()

    def get_observed_data(self):
        return self.array("rtf_ux_true").copy(), self.array("rtf_uz_true").copy()

    def set_observed_data(self, ux, uz):
        self.array("rtf_ux_true")[...] = ux
        self.array("rtf_uz_true")[...] = uz

    def set_synthetic_data(self, ux, uz):
        self.array("rtf_ux")[...] = ux
        self.array("rtf_uz")[...] = uz

    def get_kernels(self):
        return (self.array("vp_kernel").copy(), self.array("vs_kernel").copy(),
                self.array("density_v_kernel").copy())

    def get_snapshots(self):
        return tuple(self.array(n).copy() for n in self._SHAPES["accu"])

    def get_coordinates(self, in_units):
        ix, iz = np.meshgrid(np.arange(self.nx, dtype=np.float64),
                             np.arange(self.nz, dtype=np.float64), indexing="ij")
        if in_units:
            return (ix - self.np_boundary) * self.dx, (iz - self.np_boundary) * self.dz
        return ix, iz

    def get_model_vector(self):
        out = np.empty(self.free_parameters, dtype=np.float64)
        rc = self._f("get_model_vector")(self._h, out.ctypes.data_as(C.POINTER(C.c_double)), out.size)
        assert rc == 0
        return out

    def set_model_vector(self, m):
        m = np.ascontiguousarray(m, dtype=np.float64)
        rc = self._f("set_model_vector")(self._h, m.ctypes.data_as(C.POINTER(C.c_double)), m.size)
        if rc != 0:
            raise ValueError("model vector has the wrong length")

    def get_gradient_vector(self):
        out = np.empty(self.free_parameters, dtype=np.float64)
        rc = self._f("get_gradient_vector")(self._h, out.ctypes.data_as(C.POINTER(C.c_double)), out.size)
        assert rc == 0
        return out

    @property
    def misfit(self):
        return self._real("misfit")


class OracleModel(_CpuModel):
    """Our C restatement (oracle/psv_oracle.c)."""
    _prefix = "orc_"

    def __init__(self, config):
        if not os.path.exists(ORC_LIB):
            build()
        lib = _load(ORC_LIB, "orc_")
        lib.orc_create.restype = C.c_void_p
        lib.orc_create.argtypes = [C.POINTER(_OrcConfig)]
        lib.orc_check_config.restype = C.c_char_p
        lib.orc_check_config.argtypes = [C.POINTER(_OrcConfig)]
        lib.orc_calculate_l2_misfit.restype = C.c_double
        for n in ("forward", "adjoint"):
            getattr(lib, "orc_" + n).restype = C.c_int
        lib.orc_forward.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.orc_adjoint.argtypes = [C.c_void_p, C.c_int]
        lib.orc_run_model.argtypes = [C.c_void_p, C.c_int]
        for n in ("update_from_velocity", "calculate_l2_misfit", "calculate_l2_adjoint_sources",
                  "reset_kernels", "map_kernels_to_velocity"):
            getattr(lib, "orc_" + n).argtypes = [C.c_void_p]
        self._lib = lib
        cfg = config_from_ini(config) if isinstance(config, str) else dict(config)
        self.config = cfg
        which = cfg["which_source_to_fire_in_which_shot"]
        if len(cfg["ix_sources"]) != cfg["n_sources"] or len(cfg["iz_sources"]) != cfg["n_sources"] \
                or len(cfg["moment_angles"]) != cfg["n_sources"]:
            raise ValueError("Dimension mismatch between n_sources and sources.ix_sources, "
                             "sources.iz_sources or sources.moment_angles")
        if len(which) != cfg["n_shots"]:
            raise ValueError("Mismatch between n_shots and sources.which_source_to_fire_in_which_shot")
        if len(cfg["ix_receivers"]) != cfg["nr"] or len(cfg["iz_receivers"]) != cfg["nr"]:
            raise ValueError("Mismatch between nr and receivers.ix_receivers or receivers.iz_receivers")
        offs = np.zeros(cfg["n_shots"] + 1, dtype=np.int32)
        offs[1:] = np.cumsum([len(s) for s in which])
        flat = np.array([s for sub in which for s in sub], dtype=np.int32)
        keep = dict(
            ixs=np.array(cfg["ix_sources"], dtype=np.int32), izs=np.array(cfg["iz_sources"], dtype=np.int32),
            ang=np.array(cfg["moment_angles"], dtype=np.float64), offs=offs, flat=flat,
            ixr=np.array(cfg["ix_receivers"], dtype=np.int32), izr=np.array(cfg["iz_receivers"], dtype=np.int32))
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))  # noqa: E731
        c = _OrcConfig()
        for k in ("nt", "nx_inner", "nz_inner", "nx_inner_boundary", "nz_inner_boundary", "dx", "dz",
                  "dt", "np_boundary", "np_factor", "scalar_rho", "scalar_vp", "scalar_vs", "npx",
                  "npz", "peak_frequency", "source_timeshift", "delay_cycles_per_shot", "n_sources",
                  "n_shots", "nr", "snapshot_interval"):
            setattr(c, k, cfg[k])
        c.ix_sources, c.iz_sources = ip(keep["ixs"]), ip(keep["izs"])
        c.moment_angles = keep["ang"].ctypes.data_as(C.POINTER(C.c_double))
        c.shot_offsets, c.shot_sources = ip(keep["offs"]), ip(keep["flat"])
        c.ix_receivers, c.iz_receivers = ip(keep["ixr"]), ip(keep["izr"])
        err = lib.orc_check_config(C.byref(c))
        if err:
            raise ValueError(err.decode())
        self._h = lib.orc_create(C.byref(c))
        if not self._h:
            raise ValueError("orc_create failed")

    def forward_simulate(self, i_shot, store_fields=True, verbose=False, output_wavefields=False,
                         omp_threads_override=0):
        if self._lib.orc_forward(self._h, i_shot, int(store_fields)) != 0:
            raise ValueError("bad shot index")

    def adjoint_simulate(self, i_shot, verbose=False, omp_threads_override=0):
        if self._lib.orc_adjoint(self._h, i_shot) != 0:
            raise ValueError("bad shot index")

    def update_from_velocity(self):
        self._lib.orc_update_from_velocity(self._h)

    def calculate_l2_misfit(self):
        self._lib.orc_calculate_l2_misfit(self._h)

    def calculate_l2_adjoint_sources(self):
        self._lib.orc_calculate_l2_adjoint_sources(self._h)

    def reset_kernels(self):
        self._lib.orc_reset_kernels(self._h)

    def map_kernels_to_velocity(self):
        self._lib.orc_map_kernels_to_velocity(self._h)

    def run_model(self, verbose=False, simulate_adjoint=True):
        self._lib.orc_run_model(self._h, int(simulate_adjoint))


def ref_available() -> bool:
    return os.path.exists(REF_LIB)


class RefModel(_CpuModel):
    """The unmodified reference core (oracle/_ref/libpsv_ref.so via oracle/ref_shim.cpp)."""
    _prefix = "ref_"

    def __init__(self, ini_path: str, _handle=None, _lib=None):
        if _handle is not None:
            self._lib, self._h = _lib, _handle
            return
        if not ref_available():
            raise RuntimeError("oracle/_ref/libpsv_ref.so has not been built (make -C oracle)")
        lib = _load(REF_LIB, "ref_")
        lib.ref_create.restype = C.c_void_p
        lib.ref_create.argtypes = [C.c_char_p]
        lib.ref_clone.restype = C.c_void_p
        lib.ref_clone.argtypes = [C.c_void_p]
        lib.ref_last_error.restype = C.c_char_p
        lib.ref_forward.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        lib.ref_adjoint.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.ref_run_model.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.ref_load_receivers.argtypes = [C.c_void_p, C.c_int]
        lib.ref_write_kernels.argtypes = [C.c_void_p]
        lib.ref_load_model.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
        lib.ref_forward_dump.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        lib.ref_create_explicit.restype = C.c_void_p
        lib.ref_create_explicit.argtypes = [C.POINTER(FullConfig)]
        for n in ("update_from_velocity", "calculate_l2_misfit", "calculate_l2_adjoint_sources",
                  "reset_kernels", "map_kernels_to_velocity", "write_receivers", "write_sources"):
            getattr(lib, "ref_" + n).argtypes = [C.c_void_p]
        self._lib = lib
        if isinstance(ini_path, dict):  # the explicit constructor (src/fdModel.h:50-64)
            c, keep = full_config(ini_path)
            self._h = lib.ref_create_explicit(C.byref(c))
        else:
            self._h = lib.ref_create(ini_path.encode())
        if not self._h:
            raise ValueError(lib.ref_last_error().decode())

    def _chk(self, rc):
        if rc != 0:
            raise ValueError(self._lib.ref_last_error().decode())

    def copy(self):
        h = self._lib.ref_clone(self._h)
        if not h:
            raise ValueError(self._lib.ref_last_error().decode())
        return RefModel("", _handle=h, _lib=self._lib)

    def forward_simulate(self, i_shot, store_fields=True, verbose=False, output_wavefields=False,
                         omp_threads_override=0):
        if output_wavefields:
            self._chk(self._lib.ref_forward_dump(self._h, i_shot, int(store_fields), 1))
        else:
            self._chk(self._lib.ref_forward(self._h, i_shot, int(store_fields), int(verbose)))

    def adjoint_simulate(self, i_shot, verbose=False, omp_threads_override=0):
        self._chk(self._lib.ref_adjoint(self._h, i_shot, int(verbose)))

    def update_from_velocity(self):
        self._chk(self._lib.ref_update_from_velocity(self._h))

    def calculate_l2_misfit(self):
        self._chk(self._lib.ref_calculate_l2_misfit(self._h))

    def calculate_l2_adjoint_sources(self):
        self._chk(self._lib.ref_calculate_l2_adjoint_sources(self._h))

    def reset_kernels(self):
        self._chk(self._lib.ref_reset_kernels(self._h))

    def map_kernels_to_velocity(self):
        self._chk(self._lib.ref_map_kernels_to_velocity(self._h))

    def run_model(self, verbose=False, simulate_adjoint=True):
        self._chk(self._lib.ref_run_model(self._h, int(verbose), int(simulate_adjoint)))

    def write_receivers(self):
        self._chk(self._lib.ref_write_receivers(self._h))

    def load_receivers(self, verbose=False):
        self._chk(self._lib.ref_load_receivers(self._h, int(verbose)))

    def write_sources(self):
        self._chk(self._lib.ref_write_sources(self._h))

    def write_kernels(self):
        self._chk(self._lib.ref_write_kernels(self._h))

    def load_model(self, de_path, vp_path, vs_path, verbose=False):
        self._chk(self._lib.ref_load_model(self._h, de_path.encode(), vp_path.encode(), vs_path.encode(), int(verbose)))
