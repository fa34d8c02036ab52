"""ORACLE (test infrastructure): ctypes wrapper of oracle/_build/liboracle.so (oracle/c/oracle_port.c).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may import this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            subprocess.run(["make", "_build/liboracle.so"], cwd=_HERE, check=True,
                           stdout=subprocess.DEVNULL)
        _lib = C.CDLL(_SO)
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(load().oracle_num_threads())


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def contract_score(A, Bt, bitpos, col_bits, adj, node_mask, scale=1.0, inv_norm=1.0,
                   threshold=1e-5, row_lo=0, n_rows=None, want_P=False):
    """All-core C evaluation of rows [row_lo, row_lo+n_rows) x all columns.  Returns
    (sums[3], (best_x, best_p), P or None)."""
    lib = load()
    A = np.ascontiguousarray(A, dtype=np.float64)
    Bt = np.ascontiguousarray(Bt, dtype=np.float64)
    kdim = A.shape[1]
    n_cols = Bt.shape[0]
    if n_rows is None:
        n_rows = A.shape[0] - row_lo
    nbits = len(bitpos)
    bp = np.asarray(bitpos, dtype=np.int32)
    adj64 = np.zeros(64, dtype=np.uint64)
    adj64[:len(adj)] = np.asarray(adj, dtype=np.uint64)
    sums = np.zeros(3)
    bx = C.c_uint64(0)
    bpv = C.c_double(0)
    P = np.zeros(1 << nbits) if want_P else None
    lib.oracle_contract_score(
        _p(A, C.c_double), _p(Bt, C.c_double), C.c_int64(row_lo), C.c_int64(n_rows),
        C.c_int64(n_cols), C.c_int(kdim), C.c_double(scale), C.c_double(inv_norm),
        _p(bp, C.c_int), C.c_int(nbits), C.c_int(col_bits), _p(adj64, C.c_uint64),
        C.c_uint64(node_mask), C.c_double(threshold), _p(sums, C.c_double),
        C.byref(bx), C.byref(bpv), _p(P, C.c_double) if want_P else None)
    return sums, (int(bx.value), float(bpv.value)), P


def score(P, lo, adj, node_mask):
    lib = load()
    P = np.ascontiguousarray(P, dtype=np.float64)
    adj64 = np.zeros(64, dtype=np.uint64)
    adj64[:len(adj)] = np.asarray(adj, dtype=np.uint64)
    sums = np.zeros(3)
    lib.oracle_score(_p(P, C.c_double), C.c_int64(lo), C.c_int64(P.size), _p(adj64, C.c_uint64),
                     C.c_uint64(node_mask), _p(sums, C.c_double))
    return sums
