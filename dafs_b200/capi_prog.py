"""ctypes binding of dafs_cuda_guide_tree / dafs_cuda_progressive (include/dafs_cuda.h,
section 5): DAFS::build_tree and DAFS::align (src/dafs.cpp:447-495, 1518-1537).  Test /
bench side only."""
import ctypes as C

import numpy as np

from .capi import _CsrBatch, _check, lib
from .capi_pct import _as_raw

_done = False


class _Params(C.Structure):
    _fields_ = [("w", C.c_float), ("th_s", C.c_float), ("min_th_s", C.c_float), ("th_a", C.c_float),
                ("eta0", C.c_float), ("t_max", C.c_int32)]


class _Result(C.Structure):
    _fields_ = [("L", C.c_int32), ("n_rows", C.c_int32), ("seq", C.POINTER(C.c_int32)), ("bits", C.POINTER(C.c_uint8)),
                ("ss", C.POINTER(C.c_uint32)), ("score", C.c_float), ("n_nodes", C.c_int32),
                ("node_iters", C.POINTER(C.c_int32)), ("node_violated", C.POINTER(C.c_uint32)),
                ("node_score", C.POINTER(C.c_float)), ("node_L", C.POINTER(C.c_int32)), ("waves", C.c_int32),
                ("total_iters", C.c_int64), ("gpu_launches", C.c_int32), ("ms_average", C.c_float), ("ms_dd", C.c_float),
                ("ms_dd_host", C.c_float), ("ms_total", C.c_float), ("owner_", C.c_void_p)]


def _protos():
    global _done
    L = lib()
    if not _done:
        L.dafs_cuda_guide_tree.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.dafs_cuda_progressive.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(_CsrBatch), C.POINTER(_CsrBatch),
                                            C.c_void_p, C.c_void_p, C.POINTER(_Params), C.POINTER(_Result)]
        L.dafs_cuda_free_prog_result.argtypes = [C.POINTER(_Result)]
        L.dafs_cuda_free_prog_result.restype = None
        _done = True
    return L


def guide_tree(sim):
    """-> (left, right, score) over 2n-1 nodes; leaves have -1 children."""
    sim = np.ascontiguousarray(sim, np.float32)
    n = sim.shape[0]
    left = np.zeros(2 * n - 1, np.int32); right = np.zeros(2 * n - 1, np.int32); score = np.zeros(2 * n - 1, np.float32)
    _check(_protos().dafs_cuda_guide_tree(n, sim.ctypes.data, left.ctypes.data, right.ctypes.data, score.ctypes.data))
    return left, right, score


def progressive(lens, mp, bp, left, right, w=4.0, th_s=0.2, th_a=0.01, eta0=0.5, t_max=600, min_th_s=None):
    """-> dict with the root alignment (order, bits[n, L]), ss[L], score and per-node arrays."""
    L = _protos()
    lens = np.ascontiguousarray(lens, np.int32)
    n = len(lens)
    left = np.ascontiguousarray(left, np.int32); right = np.ascontiguousarray(right, np.int32)
    assert len(left) == 2 * n - 1 and len(right) == 2 * n - 1
    keep = []
    rmp, rbp = _as_raw(mp, keep), _as_raw(bp, keep)
    prm = _Params(w, th_s, th_s if min_th_s is None else min_th_s, th_a, eta0, t_max)
    res = _Result()
    _check(L.dafs_cuda_progressive(n, lens.ctypes.data_as(C.POINTER(C.c_int)), C.byref(rmp), C.byref(rbp),
                                   left.ctypes.data, right.ctypes.data, C.byref(prm), C.byref(res)))
    try:
        nn, Lc = res.n_nodes, res.L
        arr = lambda p, k, dt: np.ctypeslib.as_array(p, shape=(k,)).astype(dt, copy=True)
        out = {
            "L": Lc, "score": np.float32(res.score),
            "order": arr(res.seq, n, np.int32),
            "bits": np.ctypeslib.as_array(res.bits, shape=(n * Lc,)).reshape(n, Lc).copy(),
            "ss": arr(res.ss, Lc, np.uint32) if res.ss else np.zeros(0, np.uint32),
            "node_iters": arr(res.node_iters, nn, np.int32), "node_violated": arr(res.node_violated, nn, np.uint32),
            "node_score": arr(res.node_score, nn, np.float32), "node_L": arr(res.node_L, nn, np.int32),
            "waves": res.waves, "total_iters": int(res.total_iters), "gpu_launches": res.gpu_launches,
            "ms": {"average": res.ms_average, "dd": res.ms_dd, "dd_host": res.ms_dd_host, "total": res.ms_total},
        }
    finally:
        L.dafs_cuda_free_prog_result(C.byref(res))
    return out
