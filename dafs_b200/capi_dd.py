"""ctypes binding of the dual-decomposition part of include/dafs_cuda.h
(test / bench driver side only)."""
import ctypes as C

import numpy as np

from .capi import _check, lib

NOPAIR = 0xFFFFFFFF
_FP = C.POINTER(C.c_float)
_UP = C.POINTER(C.c_uint32)


class _DDProblem(C.Structure):
    _fields_ = [("L1", C.c_int32), ("L2", C.c_int32), ("N1", C.c_int32), ("N2", C.c_int32),
                ("p_x", _FP), ("p_y", _FP), ("p_z", _FP),
                ("w", C.c_float), ("th_s", C.c_float), ("min_th_s", C.c_float), ("th_a", C.c_float),
                ("eta0", C.c_float), ("t_max", C.c_int32)]


class _DDResult(C.Structure):
    _fields_ = [("x", _UP), ("y", _UP), ("z", _UP), ("trace", _FP), ("score", C.c_float),
                ("iters", C.c_int32), ("violated", C.c_uint32), ("status", C.c_int32), ("ncbp", C.c_int64),
                ("cycles", C.c_float * 8)]


class _DDInfo(C.Structure):
    _fields_ = [("grid", C.c_int), ("block", C.c_int), ("cols_per_lane", C.c_int), ("smem_bytes", C.c_int),
                ("ctas_per_sm", C.c_int), ("items_in_smem", C.c_uint), ("workspace_bytes", C.c_int64),
                ("arena_bytes", C.c_int64), ("total_cbp", C.c_int64), ("total_entries", C.c_int64),
                ("last_ms", C.c_float), ("packed_on_device", C.c_int)]


_proto_done = False


def _protos():
    global _proto_done
    L = lib()
    if not _proto_done:
        L.dafs_cuda_dd_plan_create.argtypes = [C.c_int, C.POINTER(_DDProblem), C.c_int, C.POINTER(C.c_void_p)]
        L.dafs_cuda_dd_plan_run.argtypes = [C.c_void_p]
        L.dafs_cuda_dd_plan_sync.argtypes = [C.c_void_p, _FP]
        L.dafs_cuda_dd_plan_fetch.argtypes = [C.c_void_p, C.POINTER(_DDResult)]
        L.dafs_cuda_dd_plan_poll.argtypes = [C.c_void_p, C.POINTER(_DDResult), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.dafs_cuda_dd_plan_info.argtypes = [C.c_void_p, C.POINTER(_DDInfo)]
        L.dafs_cuda_dd_plan_destroy.argtypes = [C.c_void_p]
        L.dafs_cuda_dd_plan_destroy.restype = None
        L.dafs_cuda_dd_solve_batch.argtypes = [C.c_int, C.POINTER(_DDProblem), C.POINTER(_DDResult)]
        L.dafs_cuda_nussinov_decode.argtypes = [C.c_int, C.c_float, C.c_float, _FP, _FP, _UP, _FP]
        L.dafs_cuda_nw_decode.argtypes = [C.c_int, C.c_int, C.c_float, _FP, _FP, _UP, _FP]
        L.dafs_cuda_nw_decode_static.argtypes = [C.c_int, C.c_int, C.c_float, _FP, _FP, _UP, _FP]
        _proto_done = True
    return L


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(_FP) if a is not None else None


class DDBatch:
    """Problems (each a dict: p_x, p_y, p_z and optional n1, n2, w, th_s, min_th_s,
    th_a, eta0, t_max) packed as dafs_dd_problem[]; keeps the arrays alive."""

    DEFAULTS = dict(n1=1, n2=1, w=4.0, th_s=0.2, th_a=0.01, eta0=0.5, t_max=600)

    def __init__(self, problems):
        self.n = len(problems)
        self.keep = []
        self.raw = (_DDProblem * max(self.n, 1))()
        self.dims = []
        for k, pr in enumerate(problems):
            d = dict(self.DEFAULTS); d.update({a: b for a, b in pr.items() if a not in ("p_x", "p_y", "p_z")})
            px, py, pz = _f32(pr["p_x"]), _f32(pr["p_y"]), _f32(pr["p_z"])
            self.keep += [px, py, pz]
            L1, L2 = pz.shape
            assert px.shape == (L1, L1) and py.shape == (L2, L2)
            r = self.raw[k]
            r.L1, r.L2, r.N1, r.N2 = L1, L2, d["n1"], d["n2"]
            r.p_x, r.p_y, r.p_z = _fp(px), _fp(py), _fp(pz)
            r.w, r.th_s, r.th_a, r.eta0, r.t_max = d["w"], d["th_s"], d["th_a"], d["eta0"], d["t_max"]
            r.min_th_s = d.get("min_th_s", d["th_s"])
            self.dims.append((L1, L2, d["t_max"]))

    def results(self, trace=False):
        res = (_DDResult * max(self.n, 1))()
        outs = []
        for k, (L1, L2, tm) in enumerate(self.dims):
            x = np.zeros(L1, np.uint32); y = np.zeros(L2, np.uint32); z = np.zeros(L1, np.uint32)
            tr = np.zeros((max(tm, 1), 3), np.float32) if trace else None
            res[k].x = x.ctypes.data_as(_UP); res[k].y = y.ctypes.data_as(_UP); res[k].z = z.ctypes.data_as(_UP)
            res[k].trace = _fp(tr)
            outs.append((x, y, z, tr))
        return res, outs

    @staticmethod
    def unpack(res, outs):
        out = []
        for k, (x, y, z, tr) in enumerate(outs):
            r = res[k]
            d = {"score": np.float32(r.score), "x": x, "y": y, "z": z, "iters": int(r.iters),
                 "violated": int(r.violated), "status": int(r.status), "ncbp": int(r.ncbp),
                 "cycles": np.array(r.cycles[:], dtype=np.float64)}
            if tr is not None:
                d["trace"] = tr[:min(int(r.iters) + 1, len(tr))]
            out.append(d)
        return out


def dd_solve_batch(problems):
    """dafs_cuda_dd_solve_batch over host matrices -> list of result dicts."""
    L = _protos()
    b = DDBatch(problems)
    res, outs = b.results()
    _check(L.dafs_cuda_dd_solve_batch(b.n, b.raw, res))
    return DDBatch.unpack(res, outs)


class DDPlan:
    """Split-phase DD call: problems resident in HBM, run() repeatable."""

    def __init__(self, problems, trace=False):
        L = _protos()
        self.batch = DDBatch(problems)
        self.trace = trace
        self._h = C.c_void_p()
        _check(L.dafs_cuda_dd_plan_create(self.batch.n, self.batch.raw, 1 if trace else 0, C.byref(self._h)))

    def run(self):
        _check(lib().dafs_cuda_dd_plan_run(self._h))

    def sync(self):
        ms = C.c_float()
        _check(lib().dafs_cuda_dd_plan_sync(self._h, C.byref(ms)))
        return float(ms.value)

    def fetch(self):
        res, outs = self.batch.results(self.trace)
        _check(lib().dafs_cuda_dd_plan_fetch(self._h, res))
        return DDBatch.unpack(res, outs)

    def poll(self, state=None):
        """Non-blocking harvest (dafs_cuda_dd_plan_poll): returns (state, fresh) where fresh maps the
        index of every problem that finished since the last poll to its result dict; pass the
        returned state to the next call."""
        if state is None:
            res, outs = self.batch.results(self.trace)
            state = {"res": res, "outs": outs, "done": (C.c_int32 * max(self.batch.n, 1))()}
        before = list(state["done"])
        n_new = C.c_int32()
        _check(lib().dafs_cuda_dd_plan_poll(self._h, state["res"], state["done"], C.byref(n_new)))
        allr = DDBatch.unpack(state["res"], state["outs"])
        fresh = {k: allr[k] for k in range(self.batch.n) if state["done"][k] and not before[k]}
        assert len(fresh) == n_new.value
        return state, fresh

    def info(self):
        i = _DDInfo()
        _check(lib().dafs_cuda_dd_plan_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in _DDInfo._fields_}

    def close(self):
        if self._h:
            lib().dafs_cuda_dd_plan_destroy(self._h)
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


def nussinov_decode(p, q=None, th=0.2, w=1.0):
    """SparseNussinov::decode on the device; q None = the plain overload."""
    L = _protos()
    p = _f32(p); q = None if q is None else _f32(q)
    ss = np.zeros(p.shape[0], np.uint32)
    s = C.c_float()
    _check(L.dafs_cuda_nussinov_decode(p.shape[0], th, w, _fp(p), _fp(q), ss.ctypes.data_as(_UP), C.byref(s)))
    return np.float32(s.value), ss


def nw_decode(p, q=None, th=0.01):
    """SparseNeedlemanWunsch initialize + decode on the device."""
    L = _protos()
    p = _f32(p); q = None if q is None else _f32(q)
    al = np.zeros(p.shape[0], np.uint32)
    s = C.c_float()
    _check(L.dafs_cuda_nw_decode(p.shape[0], p.shape[1], th, _fp(p), _fp(q), al.ctypes.data_as(_UP), C.byref(s)))
    return np.float32(s.value), al


def nw_decode_static(p, q=None, th=0.01):
    """dafs_cuda_nw_decode_static: the same decode out of the library's static device scratch
    (no cudaMalloc on the path)."""
    L = _protos()
    p = _f32(p); q = None if q is None else _f32(q)
    al = np.zeros(p.shape[0], np.uint32)
    s = C.c_float()
    _check(L.dafs_cuda_nw_decode_static(p.shape[0], p.shape[1], th, _fp(p), _fp(q), al.ctypes.data_as(_UP),
                                        C.byref(s)))
    return np.float32(s.value), al
