"""ctypes wrapper of oracle/_build/libff_oracle.so (TEST INFRASTRUCTURE ONLY --
see the header of oracle/ff_oracle.c for what it restates and what is pinned)."""
import ctypes as C
import os
import subprocess

import numpy as np

from igm_b200._abi import System, Protocol, Result, IGMK_MAX_ENVELOPES, c_f64p

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libff_oracle.so")
_lib = None


def build(force=False):
    if force or not os.path.isfile(_SO) or \
            os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "ff_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "_build/libff_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.ofr_single_point.argtypes = [C.POINTER(System), C.c_double, C.c_double, C.c_double, c_f64p, c_f64p]
        _lib.ofr_single_point.restype = C.c_int
        _lib.ofr_optimize.argtypes = [C.POINTER(System), C.POINTER(Protocol), c_f64p, C.POINTER(Result)]
        _lib.ofr_optimize.restype = C.c_int
        _lib.ofr_ranpark.argtypes = [C.c_int, C.c_int, c_f64p]
        _lib.ofr_ranpark.restype = None
    return _lib


def single_point(ps, ev_scale=1.0, env_scale=1.0, bond_prefactor=1.0):
    """-> (forces [n,3] float64, dict(total, pair, bond, env=[...]))"""
    f = np.zeros((ps.n_atoms, 3), dtype=np.float64)
    e = np.zeros(3 + IGMK_MAX_ENVELOPES, dtype=np.float64)
    lib().ofr_single_point(C.byref(ps.c), ev_scale, env_scale, bond_prefactor,
                           f.ctypes.data_as(c_f64p), e.ctypes.data_as(c_f64p))
    ne = ps.env_abck.shape[0]
    return f, {"total": e[0], "pair": e[1], "bond": e[2], "env": e[3:3 + ne].copy()}


def optimize(ps, prot):
    """-> (xyz [n,3] float64, Result)"""
    x = np.zeros((ps.n_atoms, 3), dtype=np.float64)
    r = Result()
    lib().ofr_optimize(C.byref(ps.c), C.byref(prot), x.ctypes.data_as(c_f64p), C.byref(r))
    return x, r


def ranpark(seed, n):
    out = np.zeros(n, dtype=np.float64)
    lib().ofr_ranpark(int(seed), int(n), out.ctypes.data_as(c_f64p))
    return out
