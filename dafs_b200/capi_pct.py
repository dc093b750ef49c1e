"""ctypes binding of dafs_cuda_pct (include/dafs_cuda.h, section 4): the consistency
transformation of DAFS::run() (src/dafs.cpp:1796-1827).  Test / bench side only."""
import ctypes as C

import numpy as np

from .capi import CsrBatch, DeviceCsrBatch, _CsrBatch, _check, lib

_done = False


def _protos():
    global _done
    L = lib()
    if not _done:
        L.dafs_cuda_pct.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(_CsrBatch), C.POINTER(_CsrBatch),
                                    C.c_float, C.c_float, C.c_void_p, C.POINTER(_CsrBatch), C.POINTER(_CsrBatch)]
        L.dafs_cuda_pct_device.argtypes = L.dafs_cuda_pct.argtypes
        L.dafs_cuda_pct_last_ms.argtypes = [C.POINTER(C.c_float)]
        L.dafs_cuda_pct_last_ms.restype = None
        _done = True
    return L


def _as_raw(b, keep):
    """A dafs_csr_batch over numpy arrays (or a CsrBatch the library returned)."""
    if isinstance(b, (CsrBatch, DeviceCsrBatch)):
        return b._raw
    arrs = (np.ascontiguousarray(b.row_off, np.int64), np.ascontiguousarray(b.row_ptr, np.int32),
            np.ascontiguousarray(b.nz_off, np.int64), np.ascontiguousarray(b.col, np.uint32),
            np.ascontiguousarray(b.val, np.float32))
    keep.append(arrs)
    raw = _CsrBatch()
    raw.n_units = len(arrs[0]) - 1
    raw.nnz = int(arrs[2][-1])
    raw.row_off = arrs[0].ctypes.data_as(C.POINTER(C.c_int64))
    raw.row_ptr = arrs[1].ctypes.data_as(C.POINTER(C.c_int32))
    raw.nz_off = arrs[2].ctypes.data_as(C.POINTER(C.c_int64))
    raw.col = arrs[3].ctypes.data_as(C.POINTER(C.c_uint32))
    raw.val = arrs[4].ctypes.data_as(C.POINTER(C.c_float))
    return raw


def pct(lens, mp, bp, w_pct_a=0.25, w_pct_s=0.25, device_out=False):
    """-> (sim[n, n], relaxed mp, relaxed bp).  mp / bp: a CsrBatch, a DeviceCsrBatch or any
    object with row_off / row_ptr / nz_off / col / val arrays in the dafs_csr_batch layout.
    device_out: the relaxed batches stay in HBM (DeviceCsrBatch)."""
    L = _protos()
    lens = np.ascontiguousarray(lens, np.int32)
    n = len(lens)
    keep = []
    rmp, rbp = _as_raw(mp, keep), _as_raw(bp, keep)
    sim = np.zeros((n, n), np.float32)
    omp, obp = _CsrBatch(), _CsrBatch()
    fn = L.dafs_cuda_pct_device if device_out else L.dafs_cuda_pct
    _check(fn(n, lens.ctypes.data_as(C.POINTER(C.c_int)), C.byref(rmp), C.byref(rbp),
              C.c_float(w_pct_a), C.c_float(w_pct_s), sim.ctypes.data_as(C.c_void_p),
              C.byref(omp), C.byref(obp)))
    if device_out:
        return sim, DeviceCsrBatch(omp), DeviceCsrBatch(obp)
    return sim, CsrBatch(omp, None), CsrBatch(obp, None)


def pct_last_ms():
    out = (C.c_float * 5)()
    _protos().dafs_cuda_pct_last_ms(out)
    return dict(zip(("transpose", "similarity", "relax_bp", "relax_mp", "total"), [float(v) for v in out]))
