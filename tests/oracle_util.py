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


# ---- dual decomposition: decoders and solve_by_dd -----------------------------
NOPAIR = 0xFFFFFFFF


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)




def oracle_nussinov(p, q=None, th=0.2, w=1.0):
    """SparseNussinov::decode restated (q None = the plain overload)."""
    L = oracle()
    L.oracle_nussinov.restype = C.c_float
    L.oracle_nussinov.argtypes = [C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    p = _f32(p); q = None if q is None else _f32(q)
    ss = np.zeros(p.shape[0], np.uint32)
    s = L.oracle_nussinov(p.shape[0], th, w, _fp(p), _fp(q), _fp(ss))
    return np.float32(s), ss


def ref_nussinov(p, q=None, th=0.2, w=1.0):
    L = ref()
    p = _f32(p); ss = np.zeros(p.shape[0], np.uint32)
    if q is None:
        L.ref_nussinov_plain.restype = C.c_float
        L.ref_nussinov_plain.argtypes = [C.c_int, C.c_float, C.c_void_p, C.c_void_p]
        s = L.ref_nussinov_plain(p.shape[0], th, _fp(p), _fp(ss))
    else:
        q = _f32(q)
        L.ref_nussinov_dd.restype = C.c_float
        L.ref_nussinov_dd.argtypes = [C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        s = L.ref_nussinov_dd(p.shape[0], th, w, _fp(p), _fp(q), _fp(ss))
    return np.float32(s), ss


def oracle_nw(p, q=None, th=0.01):
    """initialize + decode; returns (score, al, env[(L1+1),2], bad)."""
    L = oracle()
    L.oracle_nw_envelope.argtypes = [C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
    L.oracle_nw_decode.restype = C.c_float
    L.oracle_nw_decode.argtypes = [C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.POINTER(C.c_int)]
    p = _f32(p); q = None if q is None else _f32(q)
    L1, L2 = p.shape
    env = np.zeros((L1 + 1, 2), np.uint32)
    L.oracle_nw_envelope(L1, L2, th, _fp(p), _fp(env))
    al = np.zeros(L1, np.uint32)
    bad = C.c_int(0)
    s = L.oracle_nw_decode(L1, L2, th, _fp(p), _fp(q), _fp(env), _fp(al), C.byref(bad))
    return np.float32(s), al, env, bad.value


def ref_nw(p, q=None, th=0.01):
    L = ref()
    L.ref_nw.restype = C.c_float
    L.ref_nw.argtypes = [C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    p = _f32(p); q = None if q is None else _f32(q)
    al = np.zeros(p.shape[0], np.uint32)
    s = L.ref_nw(p.shape[0], p.shape[1], th, _fp(p), _fp(q), _fp(al))
    return np.float32(s), al


def oracle_dd(p_x, p_y, p_z, n1=1, n2=1, w=4.0, th_s=0.2, th_a=0.01, eta0=0.5, t_max=600,
              min_th_s=None, trace=False):
    L = oracle()
    L.oracle_dd_solve.restype = C.c_float
    L.oracle_dd_solve.argtypes = [C.c_int] * 4 + [C.c_void_p] * 3 + [C.c_float] * 5 + [C.c_int] + \
        [C.c_void_p] * 3 + [C.POINTER(C.c_int), C.POINTER(C.c_uint), C.POINTER(C.c_int64), C.c_void_p, C.c_int]
    p_x, p_y, p_z = _f32(p_x), _f32(p_y), _f32(p_z)
    L1, L2 = p_z.shape
    x = np.zeros(L1, np.uint32); y = np.zeros(L2, np.uint32); z = np.zeros(L1, np.uint32)
    it, vi, nc = C.c_int(0), C.c_uint(0), C.c_int64(0)
    tr = np.zeros((max(t_max, 1), 3), np.float32) if trace else None
    s = L.oracle_dd_solve(L1, L2, n1, n2, _fp(p_x), _fp(p_y), _fp(p_z), w, th_s,
                          th_s if min_th_s is None else min_th_s, th_a, eta0, t_max,
                          _fp(x), _fp(y), _fp(z), C.byref(it), C.byref(vi), C.byref(nc),
                          _fp(tr), max(t_max, 1) if trace else 0)
    out = {"score": np.float32(s), "x": x, "y": y, "z": z, "iters": it.value, "violated": vi.value,
           "ncbp": nc.value}
    if trace:
        out["trace"] = tr[:min(it.value + 1, t_max)]
    return out


def ref_dd(p_x, p_y, p_z, n1=1, n2=1, w=4.0, th_s=0.2, th_a=0.01, eta0=0.5, t_max=600):
    L = ref()
    L.ref_dd_solve.restype = C.c_float
    L.ref_dd_solve.argtypes = [C.c_int] * 4 + [C.c_void_p] * 3 + [C.c_float] * 4 + [C.c_int] + \
        [C.c_void_p] * 3 + [C.POINTER(C.c_int), C.POINTER(C.c_uint)]
    p_x, p_y, p_z = _f32(p_x), _f32(p_y), _f32(p_z)
    L1, L2 = p_z.shape
    x = np.zeros(L1, np.uint32); y = np.zeros(L2, np.uint32); z = np.zeros(L1, np.uint32)
    it, vi = C.c_int(0), C.c_uint(0)
    s = L.ref_dd_solve(L1, L2, n1, n2, _fp(p_x), _fp(p_y), _fp(p_z), w, th_s, th_a, eta0, t_max,
                       _fp(x), _fp(y), _fp(z), C.byref(it), C.byref(vi))
    return {"score": np.float32(s), "x": x, "y": y, "z": z, "iters": it.value, "violated": vi.value}


def dd_problem(length, family=0, a=0, b=1, l2=None, mutation=0.2, th=0.01, mix=False):
    """One 2-leaf DD problem as align_alignments builds it (dafs.cpp:913-940) with
    PCT and alifold off: p_x, p_y = BPP of the two sequences (synthetic_bpp), p_z =
    their ProbCons posterior (restated oracle), all dense."""
    from dafs_b200.synth import homolog_family, synthetic_bpp
    seqs = homolog_family(max(a, b) + 1, length, config_id=70 + family, mutation=mutation)
    sa, sb = seqs[a], seqs[b]
    if l2 is not None:
        sb = sb[:l2]
    p_x = synthetic_bpp(len(sa), family, a)
    p_y = synthetic_bpp(length, family, b)[:len(sb), :len(sb)].copy()
    if mix:   # conflicting stems from two other families: slow, often non-converging DD
        p_x = np.maximum(p_x, synthetic_bpp(len(sa), family + 2, 0))
        p_y = np.maximum(p_y, synthetic_bpp(length, family + 1, 1)[:len(sb), :len(sb)])
    p_z = oracle_pair_dense(sa, sb, th)
    return p_x, np.ascontiguousarray(p_y), p_z


# ---- McCaskill ------------------------------------------------------------------
BLSTAR = os.path.join(ROOT, "dafs_b200", "data", "blstar37.i32")
_mc_loaded = False


def _mc():
    global _mc_loaded
    L = oracle()
    if not _mc_loaded:
        L.oracle_mccaskill_load.argtypes = [C.c_char_p]
        assert L.oracle_mccaskill_load(BLSTAR.encode()) == 0
        L.oracle_mccaskill.restype = C.c_double
        L.oracle_mccaskill.argtypes = [C.c_char_p, C.c_int, C.c_void_p]
        L.oracle_mccaskill_bruteforce.restype = C.c_double
        L.oracle_mccaskill_bruteforce.argtypes = [C.c_char_p, C.c_int, C.c_void_p]
        L.oracle_mccaskill_csr.argtypes = [C.c_char_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_mccaskill_constrained.restype = C.c_double
        L.oracle_mccaskill_constrained.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_void_p]
        L.oracle_mccaskill_bruteforce_constrained.restype = C.c_double
        L.oracle_mccaskill_bruteforce_constrained.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_void_p]
        _mc_loaded = True
    return L


def oracle_mccaskill_flops(seq):
    """FP64 flops of the reference algorithm's sums on this sequence (instrumented restatement)."""
    L = _mc()
    L.oracle_mccaskill_count_flops.restype = C.c_double
    L.oracle_mccaskill_count_flops.argtypes = [C.c_char_p, C.c_int]
    return float(L.oracle_mccaskill_count_flops(seq, len(seq)))


def oracle_mccaskill_counted(seq):
    """(ln Z, counted flops) of one oracle_mccaskill call; safe in a thread pool (thread-local counter)."""
    L = _mc()
    L.oracle_mccaskill_last_flops.restype = C.c_double
    n = len(seq)
    P = np.zeros((n, n), np.float64)
    lnz = L.oracle_mccaskill(seq, n, P.ctypes.data)
    return float(lnz), float(L.oracle_mccaskill_last_flops())


def oracle_mccaskill_constrained(seq, con):
    """Constrained variant (RNAfold::calculate(seq, str, bp)): con uses the reference's alphabet,
    '?' free, '.' unpaired, '(' ')' a given pair."""
    n = len(seq)
    P = np.zeros((n, n), np.float64)
    lnz = _mc().oracle_mccaskill_constrained(seq, con, n, P.ctypes.data)
    return P, float(lnz)


def oracle_mccaskill_bruteforce_constrained(seq, con):
    n = len(seq)
    P = np.zeros((n, n), np.float64)
    Z = _mc().oracle_mccaskill_bruteforce_constrained(seq, con, n, P.ctypes.data)
    return P, float(Z)


def oracle_mccaskill(seq):
    """(dense FP64 P[n,n] with i<j set, ln Z) from the FP64 restatement."""
    n = len(seq)
    P = np.zeros((n, n), np.float64)
    lnz = _mc().oracle_mccaskill(seq, n, P.ctypes.data)
    return P, float(lnz)


def oracle_mccaskill_bruteforce(seq):
    n = len(seq)
    P = np.zeros((n, n), np.float64)
    Z = _mc().oracle_mccaskill_bruteforce(seq, n, P.ctypes.data)
    return P, float(Z)


def oracle_mccaskill_csr(seq, th=0.01):
    n = len(seq)
    rp = np.zeros(n + 1, np.int32); col = np.zeros(n * n // 2 + 1, np.uint32); val = np.zeros(n * n // 2 + 1, np.float32)
    nnz = _mc().oracle_mccaskill_csr(seq, n, th, rp.ctypes.data, col.ctypes.data, val.ctypes.data)
    return rp, col[:nnz].copy(), val[:nnz].copy()


# ---- consistency transformation (PCT) ---------------------------------------------
class FlatCsr:
    """A batch of row-sparse matrices in the layout of dafs_csr_batch."""

    def __init__(self, units):
        """units: list of (row_ptr[rows+1], col, val) per unit."""
        self.n_units = len(units)
        self.row_off = np.zeros(len(units) + 1, np.int64)
        self.nz_off = np.zeros(len(units) + 1, np.int64)
        for u, (rp, col, val) in enumerate(units):
            self.row_off[u + 1] = self.row_off[u] + len(rp)
            self.nz_off[u + 1] = self.nz_off[u] + len(col)
        self.row_ptr = np.concatenate([np.asarray(u[0], np.int32) for u in units]) if units else np.zeros(0, np.int32)
        self.col = np.ascontiguousarray(np.concatenate([np.asarray(u[1], np.uint32) for u in units] + [np.zeros(0, np.uint32)]))
        self.val = np.ascontiguousarray(np.concatenate([np.asarray(u[2], np.float32) for u in units] + [np.zeros(0, np.float32)]))
        self.row_ptr = np.ascontiguousarray(self.row_ptr)

    def unit(self, u):
        r0, r1 = int(self.row_off[u]), int(self.row_off[u + 1])
        z0, z1 = int(self.nz_off[u]), int(self.nz_off[u + 1])
        return self.row_ptr[r0:r1], self.col[z0:z1], self.val[z0:z1]


def pct_inputs(seqs, th_a=0.01, th_bp=0.01):
    """(lens, mp over all i<j pairs, bp per sequence) from the CPU restatements."""
    n = len(seqs)
    mp = [oracle_pair_csr(seqs[i], seqs[j], th_a)[:3] for i in range(n) for j in range(i + 1, n)]
    bp = [oracle_mccaskill_csr(s, th_bp) for s in seqs]
    return np.array([len(s) for s in seqs], np.int32), FlatCsr(mp), FlatCsr(bp)


def _pct(lib, prefix, lens, mp, bp, w_pct_a, w_pct_s):
    run, copy = getattr(lib, prefix + "_pct_run"), getattr(lib, prefix + "_pct_copy")
    run.restype = None; copy.restype = None
    run.argtypes = [C.c_int, C.c_void_p] + [C.c_void_p] * 10 + [C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    copy.argtypes = [C.c_void_p] * 6
    n = len(lens)
    lens = np.ascontiguousarray(lens, np.int32)
    sim = np.zeros((n, n), np.float32)
    sizes = np.zeros(4, np.int64)
    run(n, lens.ctypes.data, mp.row_off.ctypes.data, mp.row_ptr.ctypes.data, mp.nz_off.ctypes.data, mp.col.ctypes.data,
        mp.val.ctypes.data, bp.row_off.ctypes.data, bp.row_ptr.ctypes.data, bp.nz_off.ctypes.data, bp.col.ctypes.data,
        bp.val.ctypes.data, w_pct_a, w_pct_s, sim.ctypes.data, sizes.ctypes.data)
    mrp = np.zeros(sizes[0], np.int32); mcol = np.zeros(max(sizes[1], 1), np.uint32); mval = np.zeros(max(sizes[1], 1), np.float32)
    brp = np.zeros(sizes[2], np.int32); bcol = np.zeros(max(sizes[3], 1), np.uint32); bval = np.zeros(max(sizes[3], 1), np.float32)
    copy(mrp.ctypes.data, mcol.ctypes.data, mval.ctypes.data, brp.ctypes.data, bcol.ctypes.data, bval.ctypes.data)
    return {"sim": sim, "mp_row_ptr": mrp, "mp_col": mcol[:sizes[1]], "mp_val": mval[:sizes[1]],
            "bp_row_ptr": brp, "bp_col": bcol[:sizes[3]], "bp_val": bval[:sizes[3]]}


def oracle_pct(lens, mp, bp, w_pct_a=0.25, w_pct_s=0.25):
    return _pct(oracle(), "oracle", lens, mp, bp, w_pct_a, w_pct_s)


def ref_pct(lens, mp, bp, w_pct_a=0.25, w_pct_s=0.25):
    return _pct(ref(), "ref", lens, mp, bp, w_pct_a, w_pct_s)


# ---- progressive alignment along the guide tree --------------------------------------
def pct_to_flat(lens, r):
    """The dict _pct() returns -> (mp FlatCsr over i<j pairs, bp FlatCsr)."""
    n = len(lens)
    units, rp0, nz = [], 0, 0
    for i in range(n):
        for j in range(i + 1, n):
            rp = r["mp_row_ptr"][rp0:rp0 + lens[i] + 1]; rp0 += lens[i] + 1
            units.append((rp, r["mp_col"][nz:nz + rp[-1]], r["mp_val"][nz:nz + rp[-1]])); nz += int(rp[-1])
    bunits, rp0, nz = [], 0, 0
    for i in range(n):
        rp = r["bp_row_ptr"][rp0:rp0 + lens[i] + 1]; rp0 += lens[i] + 1
        bunits.append((rp, r["bp_col"][nz:nz + rp[-1]], r["bp_val"][nz:nz + rp[-1]])); nz += int(rp[-1])
    return FlatCsr(units), FlatCsr(bunits)


def oracle_build_tree(sim):
    n = sim.shape[0]
    sim = np.ascontiguousarray(sim, np.float32)
    left = np.zeros(2 * n - 1, np.int32); right = np.zeros(2 * n - 1, np.int32); score = np.zeros(2 * n - 1, np.float32)
    f = oracle().oracle_build_tree
    f.restype = None; f.argtypes = [C.c_int] + [C.c_void_p] * 4
    f(n, sim.ctypes.data, left.ctypes.data, right.ctypes.data, score.ctypes.data)
    return left, right, score


def _progressive(fn, is_ref, lens, mp, bp, tree=None, sim=None, w=4.0, th_s=0.2, th_a=0.01, eta0=0.5, t_max=600):
    n = len(lens)
    lens = np.ascontiguousarray(lens, np.int32)
    cap = int(lens.sum()) + 1
    nn = 2 * n - 1
    if tree is not None:
        left = np.ascontiguousarray(tree[0], np.int32).copy(); right = np.ascontiguousarray(tree[1], np.int32).copy()
        score = np.zeros(nn, np.float32) if len(tree) < 3 else np.ascontiguousarray(tree[2], np.float32).copy()
    else:
        left = np.zeros(nn, np.int32); right = np.zeros(nn, np.int32); score = np.zeros(nn, np.float32)
    L = C.c_int(0)
    order = np.zeros(n, np.int32); bits = np.zeros((n, cap), np.uint8); ss = np.zeros(cap, np.uint32)
    log = np.zeros((max(n - 1, 1), 4), np.int32); lscore = np.zeros(max(n - 1, 1), np.float32)
    csr = [mp.row_off.ctypes.data, mp.row_ptr.ctypes.data, mp.nz_off.ctypes.data, mp.col.ctypes.data, mp.val.ctypes.data,
           bp.row_off.ctypes.data, bp.row_ptr.ctypes.data, bp.nz_off.ctypes.data, bp.col.ctypes.data, bp.val.ctypes.data]
    tail = [C.c_float(w), C.c_float(th_s), C.c_float(th_a), C.c_float(eta0), C.c_int(t_max), C.c_int(cap), C.byref(L),
            C.c_void_p(order.ctypes.data), C.c_void_p(bits.ctypes.data), C.c_void_p(ss.ctypes.data),
            C.c_void_p(log.ctypes.data), C.c_void_p(lscore.ctypes.data)]
    fn.restype = C.c_float
    head = [C.c_int(n), C.c_void_p(lens.ctypes.data)] + [C.c_void_p(x) for x in csr]
    if is_ref:
        simp = None
        if tree is None:
            sim = np.ascontiguousarray(sim, np.float32); simp = C.c_void_p(sim.ctypes.data)
        s = fn(*head, simp, C.c_void_p(left.ctypes.data), C.c_void_p(right.ctypes.data), C.c_void_p(score.ctypes.data), *tail)
    else:
        s = fn(*head, C.c_void_p(left.ctypes.data), C.c_void_p(right.ctypes.data), *tail)
    Lv = L.value
    return {"score": np.float32(s), "L": Lv, "order": order, "bits": bits[:, :Lv].copy(), "ss": ss[:Lv].copy(),
            "log": log[:n - 1], "log_score": lscore[:n - 1], "left": left, "right": right, "tree_score": score}


def oracle_progressive(lens, mp, bp, tree, **kw):
    return _progressive(oracle().oracle_progressive, False, lens, mp, bp, tree=tree, **kw)


def ref_progressive(lens, mp, bp, tree=None, sim=None, **kw):
    return _progressive(ref().ref_progressive, True, lens, mp, bp, tree=tree, sim=sim, **kw)


def postorder_internal(left, right):
    """Internal nodes in the order the reference's recursion solves them (dafs.cpp:1518-1537)."""
    out = []
    def rec(ch):
        if left[ch] < 0:
            return
        rec(left[ch]); rec(right[ch]); out.append(ch)
    rec(len(left) - 1)
    return out


def stemloop_family(n, length, seed, mutation=0.12):
    """n homologs of a two-hairpin root: substitutions, short deletions and insertions (ragged
    lengths), from numpy's frozen legacy generator so every box sees the same bytes."""
    rs = np.random.RandomState(seed)
    comp = {65: 85, 67: 71, 71: 67, 85: 65}
    def hairpin(stem):
        a = rs.choice(np.frombuffer(b"ACGU", np.uint8), stem)
        return list(a) + list(b"GAAA") + [comp[int(c)] for c in a[::-1]]
    half = max(4, (length // 2 - 6) // 2)
    root = hairpin(half) + list(rs.choice(np.frombuffer(b"ACGU", np.uint8), 3)) + hairpin(max(4, (length - 2 * half - 11) // 2))
    out = []
    for _ in range(n):
        seq = []
        for c in root:
            u = rs.rand()
            if u < mutation * 0.25:
                continue                                   # deletion
            if u < mutation:
                c = int(rs.choice(np.frombuffer(b"ACGU", np.uint8)))
            seq.append(int(c))
            if rs.rand() < mutation * 0.2:
                seq.append(int(rs.choice(np.frombuffer(b"ACGU", np.uint8))))   # insertion
        out.append(bytes(seq))
    return out
