"""ctypes binding of the multi-process hand-over part of include/dafs_cuda.h (test / bench
driver side only): delivery into caller-owned host arrays (dafs_host_csr) and the NCCL
gather-to-root of per-rank fragments (dafs_comm, dafs_gathered_csr)."""
import ctypes as C

import numpy as np

from .capi import DafsCudaError, _check, _seq_args, lib

COMM_ID_BYTES = 128


class _HostCsr(C.Structure):
    _fields_ = [("cap_units", C.c_int64), ("cap_rows", C.c_int64), ("cap_nnz", C.c_int64),
                ("nz_off", C.c_void_p), ("row_ptr", C.c_void_p), ("col", C.c_void_p), ("val", C.c_void_p),
                ("n_units", C.c_int64), ("rows", C.c_int64), ("nnz", C.c_int64)]


class _Gathered(C.Structure):
    _fields_ = [("nranks", C.c_int32), ("unit_off", C.POINTER(C.c_int64)), ("nnz_off", C.POINTER(C.c_int64)),
                ("rows_off", C.POINTER(C.c_int64)), ("d_nz_off", C.c_void_p), ("d_row_ptr", C.c_void_p),
                ("d_col", C.c_void_p), ("d_val", C.c_void_p), ("owner_", C.c_void_p)]


_done = False


def _protos():
    global _done
    L = lib()
    if not _done:
        L.dafs_cuda_pairhmm_pairs_into.argtypes = [
            C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_int64,
            C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_float, C.POINTER(_HostCsr)]
        L.dafs_cuda_pairhmm_plan_fetch_into.argtypes = [C.c_void_p, C.POINTER(_HostCsr)]
        L.dafs_cuda_comm_unique_id.argtypes = [C.c_void_p]
        L.dafs_cuda_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.dafs_cuda_comm_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.dafs_cuda_comm_destroy.argtypes = [C.c_void_p]
        L.dafs_cuda_comm_destroy.restype = None
        L.dafs_cuda_pairhmm_plan_gather.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(_Gathered)]
        L.dafs_cuda_gathered_to_host.argtypes = [C.POINTER(_Gathered), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.dafs_cuda_free_gathered.argtypes = [C.POINTER(_Gathered)]
        L.dafs_cuda_free_gathered.restype = None
        _done = True
    return L


class HostCsr:
    """Caller-owned host arrays for dafs_cuda_pairhmm_pairs_into.  Pass numpy arrays (they may be
    views of shared / page-locked memory): nz_off int64[units+1], row_ptr int32[rows],
    col uint32|int32[nnz], val float32[nnz]."""

    def __init__(self, nz_off, row_ptr, col, val):
        assert nz_off.dtype == np.int64 and row_ptr.dtype == np.int32 and val.dtype == np.float32
        assert col.dtype in (np.uint32, np.int32) and len(col) == len(val)
        self.nz_off, self.row_ptr, self.col, self.val = nz_off, row_ptr, col, val
        self.raw = _HostCsr(len(nz_off) - 1, len(row_ptr), len(col), nz_off.ctypes.data, row_ptr.ctypes.data,
                            col.ctypes.data, val.ctypes.data, 0, 0, 0)

    n_units = property(lambda s: int(s.raw.n_units))
    rows = property(lambda s: int(s.raw.rows))
    nnz = property(lambda s: int(s.raw.nnz))


def pairhmm_pairs_into(seqs, pi, pj, th, dst):
    """Host sequences in, this rank's CSR out into dst (a HostCsr).  Returns dst.nnz."""
    L = _protos()
    b, n, arr, lens = _seq_args(seqs)
    pi = np.ascontiguousarray(pi, np.int32)
    pj = np.ascontiguousarray(pj, np.int32)
    _check(L.dafs_cuda_pairhmm_pairs_into(n, arr, lens, len(pi), pi.ctypes.data_as(C.POINTER(C.c_int32)),
                                          pj.ctypes.data_as(C.POINTER(C.c_int32)), C.c_float(th), C.byref(dst.raw)))
    return dst.nnz


def plan_fetch_into(plan, dst):
    _check(_protos().dafs_cuda_pairhmm_plan_fetch_into(plan._h, C.byref(dst.raw)))
    return dst.nnz


def comm_unique_id():
    buf = (C.c_ubyte * COMM_ID_BYTES)()
    _check(_protos().dafs_cuda_comm_unique_id(buf))
    return bytes(buf)


class Gathered:
    """Device-resident gather result on the root (empty elsewhere)."""

    def __init__(self, raw):
        self._raw = raw
        self.empty = not raw.owner_
        if not self.empty:
            R = raw.nranks
            self.nranks = R
            self.unit_off = np.ctypeslib.as_array(raw.unit_off, shape=(R + 1,)).copy()
            self.nnz_off = np.ctypeslib.as_array(raw.nnz_off, shape=(R + 1,)).copy()
            self.rows_off = np.ctypeslib.as_array(raw.rows_off, shape=(R + 1,)).copy()

    def to_host(self):
        """List over ranks of (nz_off, row_ptr, col, val) numpy arrays."""
        R = self.nranks
        nz = np.zeros(int(self.unit_off[R]) + R, np.int64)
        rp = np.zeros(max(int(self.rows_off[R]), 1), np.int32)
        col = np.zeros(max(int(self.nnz_off[R]), 1), np.uint32)
        val = np.zeros(max(int(self.nnz_off[R]), 1), np.float32)
        _check(_protos().dafs_cuda_gathered_to_host(C.byref(self._raw), nz.ctypes.data, rp.ctypes.data,
                                                    col.ctypes.data, val.ctypes.data))
        out = []
        for r in range(R):
            u0, u1 = int(self.unit_off[r]) + r, int(self.unit_off[r + 1]) + r + 1
            out.append((nz[u0:u1], rp[int(self.rows_off[r]):int(self.rows_off[r + 1])],
                        col[int(self.nnz_off[r]):int(self.nnz_off[r + 1])],
                        val[int(self.nnz_off[r]):int(self.nnz_off[r + 1])]))
        return out

    def close(self):
        if self._raw is not None and self._raw.owner_:
            _protos().dafs_cuda_free_gathered(C.byref(self._raw))
        self._raw = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class Comm:
    """One NCCL rank bound to the current device (dafs_comm)."""

    def __init__(self, unique_id, rank, nranks):
        self._h = C.c_void_p()
        buf = (C.c_ubyte * COMM_ID_BYTES).from_buffer_copy(unique_id)
        _check(_protos().dafs_cuda_comm_create(buf, int(rank), int(nranks), C.byref(self._h)))
        self.rank, self.nranks = rank, nranks

    def nccl_version(self):
        v = C.c_int()
        _check(_protos().dafs_cuda_comm_info(self._h, None, None, C.byref(v)))
        return int(v.value)

    def gather(self, plan, root=0):
        raw = _Gathered()
        _check(_protos().dafs_cuda_pairhmm_plan_gather(plan._h, self._h, int(root), C.byref(raw)))
        return Gathered(raw)

    def close(self):
        if self._h:
            _protos().dafs_cuda_comm_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


__all__ = ["HostCsr", "pairhmm_pairs_into", "plan_fetch_into", "comm_unique_id", "Comm", "Gathered", "DafsCudaError"]
