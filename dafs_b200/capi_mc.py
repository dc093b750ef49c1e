"""ctypes binding of the McCaskill part of include/dafs_cuda.h (test / bench side only)."""
import ctypes as C

import numpy as np

from .capi import CsrBatch, DeviceCsrBatch, _CsrBatch, _check, _seq_args, lib


class _McInfo(C.Structure):
    _fields_ = [("grid", C.c_int), ("block", C.c_int), ("smem_bytes", C.c_int), ("row_stride", C.c_int),
                ("table_bytes", C.c_int), ("workspace_bytes", C.c_int64), ("last_ms", C.c_float)]


_done = False


def _protos():
    global _done
    L = lib()
    if not _done:
        L.dafs_cuda_mccaskill_all.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_float,
                                              C.c_int, C.POINTER(_CsrBatch)]
        L.dafs_cuda_mccaskill_plan_create.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                                      C.c_float, C.POINTER(C.c_void_p)]
        L.dafs_cuda_mccaskill_plan_create_constrained.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                                                  C.POINTER(C.c_char_p), C.c_float, C.POINTER(C.c_void_p)]
        L.dafs_cuda_mccaskill_constrained.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                                      C.POINTER(C.c_char_p), C.c_float, C.POINTER(_CsrBatch)]
        L.dafs_cuda_mccaskill_plan_run.argtypes = [C.c_void_p]
        L.dafs_cuda_mccaskill_plan_sync.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int64)]
        L.dafs_cuda_mccaskill_plan_fetch.argtypes = [C.c_void_p, C.POINTER(_CsrBatch), C.c_void_p]
        L.dafs_cuda_mccaskill_plan_info.argtypes = [C.c_void_p, C.POINTER(_McInfo)]
        L.dafs_cuda_mccaskill_plan_destroy.argtypes = [C.c_void_p]
        L.dafs_cuda_mccaskill_plan_destroy.restype = None
        _done = True
    return L


def mccaskill_all(seqs, th=0.01, ngpu=1):
    """Base-pairing probabilities of every sequence -> CsrBatch (unit u = sequence u)."""
    L = _protos()
    b, n, arr, lens = _seq_args(seqs)
    raw = _CsrBatch()
    _check(L.dafs_cuda_mccaskill_all(n, arr, lens, C.c_float(th), int(ngpu), C.byref(raw)))
    return CsrBatch(raw, None)


def _cons_args(cons, n):
    """constraint strings (bytes / str / None per sequence) -> char*[n]"""
    arr = (C.c_char_p * max(n, 1))()
    keep = []
    for k, c in enumerate(cons):
        if c is not None:
            keep.append(c.encode("ascii") if isinstance(c, str) else bytes(c))
            arr[k] = keep[-1]
    return arr, keep


def mccaskill_constrained(seqs, cons, th=0.01):
    """RNAfold::calculate(seq, str, bp) for a batch: cons[k] in the reference's alphabet ('?' free,
    '.' unpaired, '(' ')' a given pair) or None."""
    L = _protos()
    b, n, arr, lens = _seq_args(seqs)
    carr, keep = _cons_args(cons, n)
    raw = _CsrBatch()
    _check(L.dafs_cuda_mccaskill_constrained(n, arr, lens, carr, C.c_float(th), C.byref(raw)))
    return CsrBatch(raw, None)


class McCaskillPlan:
    def __init__(self, seqs, th=0.01, cons=None):
        L = _protos()
        b, n, arr, lens = _seq_args(seqs)
        self.nseq = n
        self._h = C.c_void_p()
        if cons is None:
            _check(L.dafs_cuda_mccaskill_plan_create(n, arr, lens, C.c_float(th), C.byref(self._h)))
        else:
            carr, keep = _cons_args(cons, n)
            _check(L.dafs_cuda_mccaskill_plan_create_constrained(n, arr, lens, carr, C.c_float(th), C.byref(self._h)))

    def run(self):
        _check(lib().dafs_cuda_mccaskill_plan_run(self._h))

    def sync(self):
        ms, nnz = C.c_float(), C.c_int64()
        _check(lib().dafs_cuda_mccaskill_plan_sync(self._h, C.byref(ms), C.byref(nnz)))
        return float(ms.value), int(nnz.value)

    def fetch(self):
        raw = _CsrBatch()
        lnz = np.zeros(max(self.nseq, 1), np.float64)
        _check(lib().dafs_cuda_mccaskill_plan_fetch(self._h, C.byref(raw), lnz.ctypes.data_as(C.c_void_p)))
        return CsrBatch(raw, None), lnz[:self.nseq]

    def fetch_device(self):
        L = lib()
        L.dafs_cuda_mccaskill_plan_fetch_device.argtypes = [C.c_void_p, C.POINTER(_CsrBatch)]
        raw = _CsrBatch()
        _check(L.dafs_cuda_mccaskill_plan_fetch_device(self._h, C.byref(raw)))
        return DeviceCsrBatch(raw)

    def info(self):
        i = _McInfo()
        _check(lib().dafs_cuda_mccaskill_plan_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in _McInfo._fields_}

    def close(self):
        if self._h:
            lib().dafs_cuda_mccaskill_plan_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
