"""oracle/dp.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes access to the two CPU DP libraries:

* ``oracle``: our C restatement (oracle/seqalign_oracle.c);
* ``ref``:    the reference's own SequenceAlignment.c compiled in place into
  oracle/_ref/ (see oracle/build.py).  Its wrappers mirror the argument order
  of benchmarking/pySeqAlign.pyx:7-23 -> SequenceAlignment.c:73,113,147
  (qmax/dmax: rows, cols;  sw: n, m with a (n+1)x(m+1) table).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import build as _build

_u8p = ctypes.POINTER(ctypes.c_uint8)
_f32p = ctypes.POINTER(ctypes.c_float)


def _as_u8(S):
    S = np.ascontiguousarray(S, dtype=np.uint8).reshape(-1)
    return S, S.ctypes.data_as(_u8p)


def _as_f32(D):
    assert D.dtype == np.float32 and D.flags.c_contiguous
    return D.ctypes.data_as(_f32p)


class _Lib:
    def __init__(self, path, names):
        self.lib = ctypes.CDLL(path)
        self.fn = {}
        for key, sym in names.items():
            f = getattr(self.lib, sym)
            f.restype = ctypes.c_float
            f.argtypes = [_u8p, _f32p, ctypes.c_int, ctypes.c_int]
            self.fn[key] = f

    def _call(self, key, S, D, a, b):
        S, sp = _as_u8(S)
        return float(self.fn[key](sp, _as_f32(D), int(a), int(b)))

    def qmax(self, S, D, rows, cols):
        return self._call("qmax", S, D, rows, cols)

    def dmax(self, S, D, rows, cols):
        return self._call("dmax", S, D, rows, cols)

    def sw(self, S, D, n, m):
        return self._call("sw", S, D, n, m)


_oracle = None
_ref = None


def oracle_lib() -> _Lib:
    global _oracle
    if _oracle is None:
        path = _build.build_oracle()
        _oracle = _Lib(path, {"qmax": "oracle_qmax", "dmax": "oracle_dmax", "sw": "oracle_sw"})
    return _oracle


def ref_lib():
    """The compiled reference C file, or None if it is neither buildable here
    (no /root/reference) nor prebuilt in oracle/_ref/."""
    global _ref
    if _ref is None:
        path = _build.build_ref()
        if path is None or not os.path.exists(path):
            return None
        _ref = _Lib(path, {"qmax": "qmax_c", "dmax": "dmax_c", "sw": "swalignimpconstrained"})
    return _ref


def qmax_dmax_serra09(S2d, lib=None):
    """Serra09.py:173-175: one zeroed M*N table, qmax, then dmax on the SAME
    table (stale boundary).  Returns (qmax_raw, dmax_raw)."""
    lib = lib or oracle_lib()
    S2d = np.ascontiguousarray(S2d, dtype=np.uint8)
    M, N = S2d.shape
    D = np.zeros(M * N, dtype=np.float32)
    q = lib.qmax(S2d, D, M, N)
    d = lib.dmax(S2d, D, M, N)
    return q, d


def dmax_clean(S2d, lib=None):
    """dmax on a freshly zeroed table (ChenFusion.py:66 ``D *= 0``)."""
    lib = lib or oracle_lib()
    S2d = np.ascontiguousarray(S2d, dtype=np.uint8)
    M, N = S2d.shape
    D = np.zeros(M * N, dtype=np.float32)
    return lib.dmax(S2d, D, M, N)


def qmax_only(S2d, lib=None):
    lib = lib or oracle_lib()
    S2d = np.ascontiguousarray(S2d, dtype=np.uint8)
    M, N = S2d.shape
    D = np.zeros(M * N, dtype=np.float32)
    return lib.qmax(S2d, D, M, N)


def sw_score(S2d, lib=None):
    """EarlySNF_Old.py:200-201: (M+1)*(N+1) zeroed table, swconstrained(S, D, M, N)."""
    lib = lib or oracle_lib()
    S2d = np.ascontiguousarray(S2d, dtype=np.uint8)
    M, N = S2d.shape
    D = np.zeros((M + 1) * (N + 1), dtype=np.float32)
    return lib.sw(S2d, D, M, N)
