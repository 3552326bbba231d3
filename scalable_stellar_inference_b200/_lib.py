"""ctypes binding of libcsi.so (include/csi.h).  The CUDA library is mandatory: there is no CPU fallback.

``load()`` raises ``CsiLibraryError`` when the shared object is missing (run
``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C scalable_stellar_inference_b200/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcsi.so")

MAX_LAYERS, MAX_BANDS, MAX_MEAN, MAX_OBS, MAX_STAR_SITES, MAX_HYPER = 8, 4, 12, 12, 9, 5
LEVEL_SINGLE, LEVEL_MULTI, LEVEL_HIERARCHICAL = 0, 1, 2
PRIOR_FLAT, PRIOR_INFORMATIVE = 0, 1
KIND_DIAG, KIND_FULL = 0, 1
BC_NONE, BC_GRID, BC_NET = 0, 1, 2
ENGINE_TCGEN05, ENGINE_FFMA = 0, 1
OK, EINVAL, ENODEVICE, ECUDA, ESTATE = 0, -1, -2, -3, -4


class CsiLibraryError(RuntimeError):
    pass


class CsiError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libcsi error {code}: {msg}")
        self.code = code


class Net(C.Structure):
    _fields_ = [("n_layers", C.c_int32), ("n_in", C.c_int32), ("n_hidden", C.c_int32), ("n_out", C.c_int32),
                ("weights", C.POINTER(C.c_float) * MAX_LAYERS), ("bias", C.POINTER(C.c_float) * MAX_LAYERS),
                ("in_loc", C.POINTER(C.c_double)), ("in_scale", C.POINTER(C.c_double)),
                ("out_loc", C.POINTER(C.c_double)), ("out_scale", C.POINTER(C.c_double))]


class Grid(C.Structure):
    _fields_ = [("shape", C.c_int32 * 4), ("points", C.POINTER(C.c_double) * 4), ("values", C.POINTER(C.c_double))]


class Tn(C.Structure):
    _fields_ = [("loc", C.c_double), ("scale", C.c_double), ("low", C.c_double), ("high", C.c_double)]


class Config(C.Structure):
    _fields_ = [("level", C.c_int32), ("prior", C.c_int32), ("kind", C.c_int32), ("n_obs", C.c_int32),
                ("obs_index", C.c_int32 * MAX_OBS), ("bands", C.c_int32), ("bc_mode", C.c_int32),
                ("log_evol", Tn), ("log_mass", Tn), ("m_h", Tn),
                ("y_low", C.c_double), ("y_high", C.c_double), ("a_low", C.c_double), ("a_high", C.c_double),
                ("y0_loc", C.c_double), ("y0_scale", C.c_double), ("dydz", Tn), ("mu_a", Tn),
                ("py_conc", C.c_double), ("py_rate", C.c_double), ("pa_conc", C.c_double), ("pa_rate", C.c_double),
                ("delta_df", C.c_double), ("delta_loc", C.c_double * 4), ("delta_cov", C.c_double * 16),
                ("sd_c1", C.c_double), ("sd_c0", C.c_double), ("max_distance", C.c_double), ("av", C.c_double),
                ("emulator", Net), ("bc_net", Net), ("bc_grid", Grid),
                ("device", C.c_int32), ("engine", C.c_int32)]


class NutsStats(C.Structure):
    _fields_ = [("potential", C.c_double), ("accept_prob", C.c_double), ("energy0", C.c_double),
                ("num_leapfrog", C.c_int64), ("launches", C.c_int64), ("depth", C.c_int32), ("diverging", C.c_int32),
                ("turning", C.c_int32), ("host_syncs", C.c_int32)]


PEER_SLOT_DOUBLES = 32
UTURN_STATIC_MAX_LEN = 1 << 24


_PROTOS = {
    "csi_last_error": (C.c_char_p, []),
    "csi_version": (C.c_int, []),
    "csi_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "csi_destroy": (None, [C.c_void_p]),
    "csi_set_shard": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "csi_set_peer_exchange": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "csi_set_data": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "csi_set_grid_limit": (C.c_int, [C.c_void_p, C.c_int32]),
    "csi_layout": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_char_p)]),
    "csi_potential": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "csi_potential_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "csi_leapfrog_pre": (C.c_int, [C.c_void_p] * 5 + [C.c_float] + [C.c_void_p] * 4),
    "csi_leapfrog_post": (C.c_int, [C.c_void_p] * 5 + [C.c_float] + [C.c_void_p] * 4),
    "csi_uturn": (C.c_int, [C.c_void_p] * 7),
    "csi_uturn_static": (C.c_int, [C.c_void_p] * 7),
    "csi_nuts_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_double, C.c_uint64,
                                  C.c_int64, C.c_int64, C.POINTER(C.c_void_p)]),
    "csi_nuts_destroy": (None, [C.c_void_p]),
    "csi_nuts_init_state": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "csi_nuts_transition": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64, C.POINTER(NutsStats)]),
    "csi_nuts_probe_step": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64, C.POINTER(C.c_double)]),
    "csi_nuts_begin_only": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64]),
    "csi_nuts_peek": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p]),
    "csi_nuts_hash": (C.c_uint64, [C.c_uint64] * 4),
    "csi_net_forward": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "csi_net_vjp": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "csi_grid_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "csi_grid_vjp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "csi_predict": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "csi_dev_alloc": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]),
    "csi_dev_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "csi_memcpy_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "csi_memcpy_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "csi_synchronize": (C.c_int, [C.c_void_p]),
    "csi_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "csi_time_potential": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                     C.POINTER(C.c_float)]),
}

EXPORTS = tuple(_PROTOS)
_lib = None


def build(verbose=False):
    """Compile libcsi.so (and the hardware probe) for sm_100a with nvcc; needs no GPU."""
    out = subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if verbose or out.returncode:
        print(out.stdout[-4000:], out.stderr[-4000:])
    if out.returncode:
        raise CsiLibraryError("building libcsi.so failed")
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CsiLibraryError(f"{LIB_PATH} is missing: build it with `make -C {os.path.join(HERE, 'csrc')}`; "
                              "there is no CPU fallback for the potential")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc):
    if rc != OK:
        raise CsiError(rc, load().csi_last_error().decode())
