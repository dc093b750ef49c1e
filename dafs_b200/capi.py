"""ctypes binding of include/dafs_cuda.h (test / bench driver side only)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def library_path():
    """In-tree libdafs_cuda.so; DAFS_CUDA_LIB points at another build (tuning experiments)."""
    return os.environ.get("DAFS_CUDA_LIB") or os.path.join(_HERE, "libdafs_cuda.so")


class DafsCudaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libdafs_cuda error %d: %s" % (code, msg))
        self.code = code


class _CsrBatch(C.Structure):
    _fields_ = [
        ("n_units", C.c_int64),
        ("nnz", C.c_int64),
        ("row_off", C.POINTER(C.c_int64)),
        ("row_ptr", C.POINTER(C.c_int32)),
        ("nz_off", C.POINTER(C.c_int64)),
        ("col", C.POINTER(C.c_uint32)),
        ("val", C.POINTER(C.c_float)),
        ("owner_", C.c_void_p),
    ]


class _PairHmmInfo(C.Structure):
    _fields_ = [("cols_per_lane", C.c_int), ("grid", C.c_int), ("warps_per_cta", C.c_int),
                ("launches_per_run", C.c_int), ("workspace_bytes", C.c_int64),
                ("last_total_ms", C.c_float), ("last_kernel_ms", C.c_float)]


_lib = None


def lib():
    """Load libdafs_cuda.so; fail loudly when it is not built."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise ImportError(
                "%s is missing: build it with `make` (or __graft_entry__.build()); "
                "there is no CPU fallback" % path)
        L = C.CDLL(path)
        L.dafs_cuda_last_error.restype = C.c_char_p
        L.dafs_cuda_pairhmm_plan_create.argtypes = [
            C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_int64,
            C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_float, C.POINTER(C.c_void_p)]
        L.dafs_cuda_pairhmm_plan_run.argtypes = [C.c_void_p]
        L.dafs_cuda_pairhmm_plan_sync.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int64)]
        L.dafs_cuda_pairhmm_plan_fetch.argtypes = [C.c_void_p, C.POINTER(_CsrBatch)]
        L.dafs_cuda_pairhmm_plan_destroy.argtypes = [C.c_void_p]
        L.dafs_cuda_pairhmm_plan_export.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_int64]
        L.dafs_cuda_pairhmm_plan_info.argtypes = [C.c_void_p, C.POINTER(_PairHmmInfo)]
        L.dafs_cuda_pairhmm_plan_destroy.restype = None
        L.dafs_cuda_pairhmm_pairs.argtypes = [
            C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_int64,
            C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_float, C.POINTER(_CsrBatch)]
        L.dafs_cuda_pairhmm_all.argtypes = [
            C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_float, C.c_int,
            C.POINTER(_CsrBatch)]
        L.dafs_cuda_free_batch.argtypes = [C.POINTER(_CsrBatch)]
        L.dafs_cuda_free_batch.restype = None
        L.dafs_cuda_pairhmm_tables.argtypes = [C.POINTER(C.c_float)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise DafsCudaError(rc, lib().dafs_cuda_last_error().decode("utf-8", "replace"))


def init(ngpu=0):
    _check(lib().dafs_cuda_init(int(ngpu)))


def device_count():
    return int(lib().dafs_cuda_device_count())


def fp32_probe():
    """(FFMA TFLOP/s, FMUL+FADD TFLOP/s) measured on the current device."""
    a, b = C.c_double(0), C.c_double(0)
    L = lib()
    L.dafs_cuda_fp32_probe.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double)]
    _check(L.dafs_cuda_fp32_probe(C.byref(a), C.byref(b)))
    return a.value, b.value


def fp64_probe():
    """DFMA TFLOP/s measured on the current device."""
    a = C.c_double(0)
    L = lib()
    L.dafs_cuda_fp64_probe.argtypes = [C.POINTER(C.c_double)]
    _check(L.dafs_cuda_fp64_probe(C.byref(a)))
    return a.value


def _as_bytes(seqs):
    return [s.encode("ascii") if isinstance(s, str) else bytes(s) for s in seqs]


def _seq_args(seqs):
    b = _as_bytes(seqs)
    n = len(b)
    arr = (C.c_char_p * max(n, 1))(*b)
    lens = (C.c_int * max(n, 1))(*[len(x) for x in b])
    return b, n, arr, lens


def all_pairs(n):
    """(pi, pj) of every i<j in the reference's order (src/align.cpp:41-51)."""
    pi, pj = np.triu_indices(n, k=1)
    return pi.astype(np.int32), pj.astype(np.int32)


class DeviceCsrBatch:
    """Owns a device-resident dafs_csr_batch (row_off / nz_off on the host, row_ptr / col / val in
    HBM): pass it to pct / progressive, or make a host copy with to_host()."""

    def __init__(self, raw):
        self._raw = raw
        n = int(raw.n_units)
        self.n_units, self.nnz = n, int(raw.nnz)
        self.row_off = np.ctypeslib.as_array(raw.row_off, shape=(n + 1,))
        self.nz_off = np.ctypeslib.as_array(raw.nz_off, shape=(n + 1,))

    def to_host(self):
        L = lib()
        L.dafs_cuda_batch_to_host.argtypes = [C.POINTER(_CsrBatch), C.POINTER(_CsrBatch)]
        out = _CsrBatch()
        _check(L.dafs_cuda_batch_to_host(C.byref(self._raw), C.byref(out)))
        return CsrBatch(out, None)

    def close(self):
        if self._raw is not None:
            lib().dafs_cuda_free_batch(C.byref(self._raw))
            self._raw = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CsrBatch:
    """Owns a dafs_csr_batch; numpy views are valid until close()."""

    def __init__(self, raw, rows_of_unit):
        self._raw = raw
        n = int(raw.n_units)
        self.n_units = n
        self.nnz = int(raw.nnz)
        self.row_off = np.ctypeslib.as_array(raw.row_off, shape=(n + 1,)) if n >= 0 else None
        self.nz_off = np.ctypeslib.as_array(raw.nz_off, shape=(n + 1,))
        total_rows = int(self.row_off[n])
        self.row_ptr = (np.ctypeslib.as_array(raw.row_ptr, shape=(total_rows,))
                        if total_rows else np.zeros(0, np.int32))
        self.col = np.ctypeslib.as_array(raw.col, shape=(self.nnz,)) if self.nnz else np.zeros(0, np.uint32)
        self.val = np.ctypeslib.as_array(raw.val, shape=(self.nnz,)) if self.nnz else np.zeros(0, np.float32)

    def unit(self, u):
        """(row_ptr[rows+1], col[nnz_u], val[nnz_u]) of unit u (views)."""
        r0, r1 = int(self.row_off[u]), int(self.row_off[u + 1])
        z0, z1 = int(self.nz_off[u]), int(self.nz_off[u + 1])
        return self.row_ptr[r0:r1], self.col[z0:z1], self.val[z0:z1]

    def dense(self, u, ncols):
        rp, col, val = self.unit(u)
        rows = len(rp) - 1
        d = np.zeros((rows, ncols), np.float32)
        rr = np.repeat(np.arange(rows), np.diff(rp))
        d[rr, col] = val
        return d

    def close(self):
        if self._raw is not None:
            lib().dafs_cuda_free_batch(C.byref(self._raw))
            self._raw = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def pairhmm_tables():
    out = (C.c_float * 54)()
    _check(lib().dafs_cuda_pairhmm_tables(out))
    return np.array(out, dtype=np.float32)


def pairhmm_pairs(seqs, pi, pj, th=0.01):
    b, n, arr, lens = _seq_args(seqs)
    pi = np.ascontiguousarray(pi, np.int32)
    pj = np.ascontiguousarray(pj, np.int32)
    raw = _CsrBatch()
    _check(lib().dafs_cuda_pairhmm_pairs(
        n, arr, lens, len(pi), pi.ctypes.data_as(C.POINTER(C.c_int32)),
        pj.ctypes.data_as(C.POINTER(C.c_int32)), C.c_float(th), C.byref(raw)))
    return CsrBatch(raw, None)


def pairhmm_all(seqs, th=0.01, ngpu=1):
    b, n, arr, lens = _seq_args(seqs)
    raw = _CsrBatch()
    _check(lib().dafs_cuda_pairhmm_all(n, arr, lens, C.c_float(th), int(ngpu), C.byref(raw)))
    return CsrBatch(raw, None)


class PairHmmPlan:
    """Split-phase pair-HMM call: inputs resident in HBM between run()s."""

    def __init__(self, seqs, pi, pj, th=0.01):
        b, n, arr, lens = _seq_args(seqs)
        pi = np.ascontiguousarray(pi, np.int32)
        pj = np.ascontiguousarray(pj, np.int32)
        self._h = C.c_void_p()
        _check(lib().dafs_cuda_pairhmm_plan_create(
            n, arr, lens, len(pi), pi.ctypes.data_as(C.POINTER(C.c_int32)),
            pj.ctypes.data_as(C.POINTER(C.c_int32)), C.c_float(th), C.byref(self._h)))
        self.npairs = len(pi)

    def run(self):
        _check(lib().dafs_cuda_pairhmm_plan_run(self._h))

    def sync(self):
        ms = C.c_float()
        nnz = C.c_int64()
        _check(lib().dafs_cuda_pairhmm_plan_sync(self._h, C.byref(ms), C.byref(nnz)))
        return float(ms.value), int(nnz.value)

    def info(self):
        i = _PairHmmInfo()
        _check(lib().dafs_cuda_pairhmm_plan_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in _PairHmmInfo._fields_}

    def export(self, d_row_ptr, d_nz_off, d_col, d_val, cap_nnz):
        """Raw device pointers (ints) of caller-owned buffers on this plan's device."""
        _check(lib().dafs_cuda_pairhmm_plan_export(self._h, d_row_ptr, d_nz_off, d_col, d_val,
                                                   int(cap_nnz)))

    def fetch(self):
        raw = _CsrBatch()
        _check(lib().dafs_cuda_pairhmm_plan_fetch(self._h, C.byref(raw)))
        return CsrBatch(raw, None)

    def fetch_device(self):
        """The result as a device-resident batch (the plan cannot be run again)."""
        L = lib()
        L.dafs_cuda_pairhmm_plan_fetch_device.argtypes = [C.c_void_p, C.POINTER(_CsrBatch)]
        raw = _CsrBatch()
        _check(L.dafs_cuda_pairhmm_plan_fetch_device(self._h, C.byref(raw)))
        return DeviceCsrBatch(raw)

    def close(self):
        if self._h:
            lib().dafs_cuda_pairhmm_plan_destroy(self._h)
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
