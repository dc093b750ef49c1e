"""ctypes access to the checkers: oracle/libdafs_oracle.so (this repo's C
restatement) and, when built, oracle/_ref/libdafs_ref.so (the reference's own
sources).  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libdafs_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libdafs_ref.so")

_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), ORACLE_SO])
        L = C.CDLL(ORACLE_SO)
        L.oracle_pairhmm_pair.restype = C.c_int
        L.oracle_pairhmm_pair.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_float,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_pairhmm_all.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                         C.c_float, C.c_long, C.c_long, C.c_int, C.c_void_p]
        _oracle = L
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        L.ref_probcons_pair.restype = C.c_int
        L.ref_probcons_pair.argtypes = [C.c_char_p, C.c_char_p, C.c_float, C.c_void_p, C.c_void_p]
        L.ref_probcons_all.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.c_float, C.c_long, C.c_long,
                                       C.c_int, C.c_void_p]
        _ref = L
    return _ref


def oracle_pair_csr(s1, s2, th=0.01):
    """(row_ptr, col, val, total) from the C restatement."""
    L1, L2 = len(s1), len(s2)
    row_ptr = np.zeros(L1 + 1, np.int32)
    col = np.zeros(L1 * L2, np.uint32)
    val = np.zeros(L1 * L2, np.float32)
    dbg = np.zeros(4, np.float32)
    n = oracle().oracle_pairhmm_pair(s1, L1, s2, L2, th, None, row_ptr.ctypes.data,
                                     col.ctypes.data, val.ctypes.data, dbg.ctypes.data)
    assert n >= 0
    return row_ptr, col[:n].copy(), val[:n].copy(), float(dbg[0])


def oracle_pair_dense(s1, s2, th=0.01):
    L1, L2 = len(s1), len(s2)
    d = np.zeros((L1, L2), np.float32)
    n = oracle().oracle_pairhmm_pair(s1, L1, s2, L2, th, d.ctypes.data, None, None, None, None)
    assert n >= 0
    return d


def ref_pair_dense(s1, s2, th=0.01):
    """(dense, similarity score) from the reference's ProbCons::calculate."""
    L1, L2 = len(s1), len(s2)
    d = np.zeros((L1, L2), np.float32)
    sim = C.c_float()
    ref().ref_probcons_pair(s1, s2, th, d.ctypes.data, C.addressof(sim))
    return d, float(sim.value)
