"""TEST INFRASTRUCTURE — ctypes driver for the C restatements in oracle/ (liboracle.so):
DCT-I (dct1.c), fourierSolve + gradient merge (poisson_oracle.c), PCG32 (pcg_oracle.c).
Only tests/, __graft_entry__.smoke() and bench.py's CPU arm may import this."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None
dp = C.POINTER(C.c_double)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.gdpt_oracle_dct1_direct.argtypes = [dp, dp, C.c_int]
        L.gdpt_oracle_dct1_1d.argtypes = [dp, C.c_int]
        L.gdpt_oracle_dct1_2d.argtypes = [dp, C.c_int, C.c_int]
        L.gdpt_oracle_poisson.argtypes = [C.c_int, C.c_int, dp, dp, dp, C.c_double, dp]
        L.gdpt_oracle_merge.argtypes = [C.c_int, C.c_int, dp, dp, dp, dp, dp, dp]
        L.gdpt_oracle_pcg_fill.argtypes = [C.c_uint64, C.c_uint64, C.c_int, dp, C.POINTER(C.c_uint32)]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(dp)


def dct1_direct(x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.zeros_like(x)
    lib().gdpt_oracle_dct1_direct(_p(x), _p(out), x.size)
    return out


def dct1_1d(x):
    x = np.array(x, dtype=np.float64)
    lib().gdpt_oracle_dct1_1d(_p(x), x.size)
    return x


def dct1_2d(a):
    a = np.array(a, dtype=np.float64)
    lib().gdpt_oracle_dct1_2d(_p(a), a.shape[0], a.shape[1])
    return a


def poisson(u, gx, gy, alpha=0.04):
    u, gx, gy = (np.ascontiguousarray(v, dtype=np.float64) for v in (u, gx, gy))
    h, w, _ = u.shape
    out = np.zeros_like(u)
    lib().gdpt_oracle_poisson(w, h, _p(u), _p(gx), _p(gy), alpha, _p(out))
    return out


def merge(cx0, cx1, cy0, cy1):
    cx0, cx1, cy0, cy1 = (np.ascontiguousarray(v, dtype=np.float64) for v in (cx0, cx1, cy0, cy1))
    h, w, _ = cx0.shape
    cx, cy = np.zeros_like(cx0), np.zeros_like(cx0)
    lib().gdpt_oracle_merge(w, h, _p(cx0), _p(cx1), _p(cy0), _p(cy1), _p(cx), _p(cy))
    return cx, cy


def pcg_fill(stream, n, seed=0x31E241F862A1FB5E):
    out = np.zeros(n)
    raw = np.zeros(n, dtype=np.uint32)
    lib().gdpt_oracle_pcg_fill(stream, seed, n, _p(out), raw.ctypes.data_as(C.POINTER(C.c_uint32)))
    return out, raw
