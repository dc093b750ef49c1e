"""CPU tests: oracle/dd_oracle.c (the C restatement of SparseNussinov,
SparseNeedlemanWunsch and DAFS::solve_by_dd) bit-for-bit against oracle/_ref
(the reference's own translation units) and against tests/golden/dd_golden.json,
which tests/golden/make_dd_golden.py generated from oracle/_ref."""
import json
import os
import zlib

import numpy as np
import pytest

import oracle_util as ou
from dafs_b200.synth import synthetic_bpp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
need_ref = pytest.mark.skipif(not ou.have_ref(), reason="oracle/_ref not built (no /root/reference here)")


def rand_q(shape, seed, density=0.05, scale=0.5, signed=True):
    r = np.random.default_rng(seed)
    q = (r.random(shape) < density) * (r.random(shape) - (0.5 if signed else 0.0)) * scale
    return q.astype(np.float32)


@need_ref
@pytest.mark.parametrize("L,seed", [(1, 0), (2, 1), (3, 2), (4, 3), (5, 4), (17, 5), (73, 6), (150, 7), (301, 8)])
def test_nussinov_vs_reference(L, seed):
    p = synthetic_bpp(L, seed, 0)
    q = rand_q((L, L), seed)
    for w in (4.0, 2.6666667, 1.0):
        so, sso = ou.oracle_nussinov(p, q, 0.2, w)
        sr, ssr = ou.ref_nussinov(p, q, 0.2, w)
        assert so.view(np.uint32) == sr.view(np.uint32)
        assert np.array_equal(sso, ssr)
    so, sso = ou.oracle_nussinov(p, None, 0.15)
    sr, ssr = ou.ref_nussinov(p, None, 0.15)
    assert so.view(np.uint32) == sr.view(np.uint32) and np.array_equal(sso, ssr)


@need_ref
@pytest.mark.parametrize("seed", range(6))
def test_nussinov_dense_random_vs_reference(seed):
    """Dense random p with many ties (quantised values) stresses the tie-break order."""
    r = np.random.default_rng(100 + seed)
    L = 40 + 7 * seed
    p = (np.round(r.random((L, L)) * 8) / 8 * (r.random((L, L)) < 0.3)).astype(np.float32)
    q = (np.round(r.random((L, L)) * 4) / 8 * (r.random((L, L)) < 0.2)).astype(np.float32)
    so, sso = ou.oracle_nussinov(p, q, 0.25, 1.0)
    sr, ssr = ou.ref_nussinov(p, q, 0.25, 1.0)
    assert so.view(np.uint32) == sr.view(np.uint32) and np.array_equal(sso, ssr)


@need_ref
@pytest.mark.parametrize("L1,L2,seed", [(1, 1, 0), (1, 9, 1), (9, 1, 2), (40, 40, 3), (73, 80, 4), (150, 120, 5), (300, 300, 6)])
def test_nw_vs_reference(L1, L2, seed):
    _, _, p = ou.dd_problem(max(L1, L2), family=seed, l2=None)
    p = np.ascontiguousarray(p[:L1, :L2])
    q = rand_q((L1, L2), seed, signed=False)
    for qq in (q, None):
        so, alo, _, bad = ou.oracle_nw(p, qq, 0.01)
        sr, alr = ou.ref_nw(p, qq, 0.01)
        assert bad == 0
        assert so.view(np.uint32) == sr.view(np.uint32) and np.array_equal(alo, alr)


@need_ref
@pytest.mark.parametrize("seed", range(4))
def test_nw_sparse_envelope_vs_reference(seed):
    """Rows without any alignable point and a ragged envelope."""
    r = np.random.default_rng(seed)
    L1, L2 = 30 + seed, 45 - seed
    p = ((r.random((L1, L2)) < 0.04) * r.random((L1, L2))).astype(np.float32)
    so, alo, _, bad = ou.oracle_nw(p, None, 0.3)
    sr, alr = ou.ref_nw(p, None, 0.3)
    assert bad == 0 and so.view(np.uint32) == sr.view(np.uint32) and np.array_equal(alo, alr)


@need_ref
@pytest.mark.parametrize("L,fam,n1,n2,tmax", [(30, 0, 1, 1, 600), (73, 1, 1, 1, 600), (73, 2, 3, 2, 600),
                                              (120, 3, 1, 4, 100), (150, 4, 1, 1, 60)])
def test_dd_vs_reference(L, fam, n1, n2, tmax):
    p_x, p_y, p_z = ou.dd_problem(L, family=fam)
    o = ou.oracle_dd(p_x, p_y, p_z, n1, n2, t_max=tmax)
    r = ou.ref_dd(p_x, p_y, p_z, n1, n2, t_max=tmax)
    assert o["score"].view(np.uint32) == r["score"].view(np.uint32)
    assert o["iters"] == r["iters"] and o["violated"] == r["violated"]
    for k in "xyz":
        assert np.array_equal(o[k], r[k])


@need_ref
def test_dd_empty_cbp_nan_path_vs_reference():
    """No consensus base pair: eta becomes NaN through 0.0f/0.0 (dafs.cpp:1285)."""
    p_x, p_y, p_z = ou.dd_problem(40, family=9)
    p_z[:] = 0
    o = ou.oracle_dd(p_x, p_y, p_z, t_max=20)
    r = ou.ref_dd(p_x, p_y, p_z, t_max=20)
    assert o["ncbp"] == 0
    assert o["score"].view(np.uint32) == r["score"].view(np.uint32)
    assert o["iters"] == r["iters"] and o["violated"] == r["violated"]
    for k in "xyz":
        assert np.array_equal(o[k], r[k])


GOLD_PATH = os.path.join(ROOT, "tests", "golden", "dd_golden.json")


def _gold():
    return json.load(open(GOLD_PATH))["cases"] if os.path.exists(GOLD_PATH) else []


@pytest.mark.parametrize("case", _gold(), ids=[c["name"] for c in _gold()])
def test_dd_oracle_matches_reference_golden(case):
    p_x, p_y, p_z = ou.dd_problem(case["L"], family=case["family"], l2=case.get("l2"),
                                    mutation=case["mutation"], mix=case["mix"])
    o = ou.oracle_dd(p_x, p_y, p_z, case["n1"], case["n2"], t_max=case["t_max"])
    assert int(o["score"].view(np.uint32)) == case["score_bits"]
    assert o["iters"] == case["iters"] and o["violated"] == case["violated"]
    for k in "xyz":
        assert zlib.crc32(o[k].tobytes()) == case["crc_" + k]
