"""GPU: dafs_cuda_progressive (DAFS::align along the guide tree, src/dafs.cpp:1518-1537, with
average_matching_probability / average_basepairing_probability / solve_by_dd /
project_alignment per node) through the C ABI, bit for bit against oracle/prog_oracle.c
(itself pinned against the reference's own functions, tests/test_oracle_progressive.py) and
the committed golden fixture."""
import json
import os

import numpy as np
import pytest

import oracle_util as ou
import dafs_b200 as db
from dafs_b200 import random_rna

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def inputs(seqs, w_pct=0.25):
    lens, mp, bp = ou.pct_inputs(seqs)
    r = ou.oracle_pct(lens, mp, bp, w_pct, w_pct)
    mp2, bp2 = ou.pct_to_flat(lens, r)
    return lens, mp2, bp2, r["sim"]


def check(lens, mp, bp, left, right, **kw):
    o = ou.oracle_progressive(lens, mp, bp, (left, right), **kw)
    g = db.progressive(lens, mp, bp, left, right, **kw)
    assert g["L"] == o["L"]
    assert np.array_equal(g["order"], o["order"])
    assert np.array_equal(g["bits"], o["bits"])
    assert np.array_equal(g["ss"], o["ss"])
    assert g["score"].view(np.uint32) == o["score"].view(np.uint32)
    post = ou.postorder_internal(left, right)
    assert np.array_equal(g["node_iters"][post], o["log"][:, 2])
    assert np.array_equal(g["node_violated"][post].astype(np.int64), o["log"][:, 3].astype(np.int64))
    assert np.array_equal(g["node_score"][post].view(np.uint32), o["log_score"].view(np.uint32))
    assert g["gpu_launches"] >= 3 * g["waves"]
    return g


@pytest.mark.parametrize("n,length,seed", [(2, 40, 1), (3, 50, 2), (5, 60, 3), (8, 45, 4), (6, 70, 5), (12, 90, 6)])
def test_family(n, length, seed):
    seqs = ou.stemloop_family(n, length, seed)
    lens, mp, bp, sim = inputs(seqs)
    left, right, _ = db.guide_tree(sim)
    g = check(lens, mp, bp, left, right)
    assert g["waves"] >= 1 and g["node_L"][-1] == g["L"]


def test_schedules_agree():
    """Dependency-driven scheduling (default: a node starts as soon as both children are solved,
    single problems harvested with dafs_cuda_dd_plan_poll) and the level-synchronous order
    (DAFS_PROG_WAVES=1) give the same alignment, structure, scores and per-node iterations."""
    seqs = ou.stemloop_family(14, 80, 21, mutation=0.35)
    lens, mp, bp, sim = inputs(seqs)
    left, right, _ = db.guide_tree(sim)
    a = check(lens, mp, bp, left, right, t_max=150)
    old = os.environ.get("DAFS_PROG_WAVES")
    os.environ["DAFS_PROG_WAVES"] = "1"
    try:
        b = check(lens, mp, bp, left, right, t_max=150)
    finally:
        if old is None:
            del os.environ["DAFS_PROG_WAVES"]
        else:
            os.environ["DAFS_PROG_WAVES"] = old
    for k in ("order", "bits", "ss", "node_iters", "node_violated", "node_L"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["node_score"].view(np.uint32), b["node_score"].view(np.uint32))
    assert a["total_iters"] == b["total_iters"]
    assert b["waves"] <= a["waves"]              # levels vs batches


def test_chain_tree_unrelated():
    # a chain-shaped tree over unrelated sequences: mp_[a][b] with a > b (transposes), long runs
    seqs = random_rna(5, 48, config_id=77)
    lens, mp, bp, _ = inputs(seqs)
    n = len(seqs)
    left = np.full(2 * n - 1, -1, np.int32); right = np.full(2 * n - 1, -1, np.int32)
    left[n], right[n] = 3, 1
    for k in range(n + 1, 2 * n - 1):
        left[k], right[k] = [0, 2, 4][k - n - 1], k - 1
    g = check(lens, mp, bp, left, right, t_max=200)
    assert g["waves"] == n - 1


def test_raw_posteriors_and_options():
    # no consistency transformation (-p 0 -q 0), other weights / thresholds
    seqs = ou.stemloop_family(6, 56, 9)
    lens, mp, bp = ou.pct_inputs(seqs)
    r = ou.oracle_pct(lens, mp, bp, 0.0, 0.0)
    left, right, _ = db.guide_tree(r["sim"])
    check(lens, mp, bp, left, right, w=2.0, th_s=0.3, th_a=0.05, eta0=0.25, t_max=100)


def test_single_sequence():
    seqs = random_rna(1, 30, config_id=5)
    lens, mp, bp = ou.pct_inputs(seqs)
    g = db.progressive(lens, mp, bp, [-1], [-1])
    assert g["L"] == 30 and g["bits"].all() and g["score"] == 0.0 and g["waves"] == 0


def test_golden():
    g = json.load(open(os.path.join(HERE, "golden", "prog_golden.json")))
    for case in g["cases"]:
        seqs = [s.encode() for s in case["seqs"]]
        lens, mp, bp, sim = inputs(seqs)
        left, right, _ = db.guide_tree(sim)
        assert left.tolist() == case["left"] and right.tolist() == case["right"]
        r = db.progressive(lens, mp, bp, left, right)
        assert r["L"] == case["L"] and r["order"].tolist() == case["order"]
        assert ["".join("1" if b else "0" for b in row) for row in r["bits"]] == case["bits"]
        assert [int(v) for v in r["ss"].view(np.int32)] == case["ss"]
        assert int(r["score"].view(np.uint32)) == case["score_bits"]
        post = ou.postorder_internal(left, right)
        assert r["node_iters"][post].tolist() == [row[2] for row in case["log"]]
        assert r["node_score"][post].view(np.uint32).tolist() == case["log_score_bits"]


def test_full_device_pipeline_matches_oracle_pipeline():
    # posteriors, PCT, tree and progressive alignment all through the C ABI
    seqs = ou.stemloop_family(7, 64, 21)
    lens = np.array([len(s) for s in seqs], np.int32)
    with db.pairhmm_all(seqs, 0.01) as mp, db.mccaskill_all(seqs, 0.01) as bp:
        sim, mp2, bp2 = db.pct(lens, mp, bp, 0.25, 0.25)
        with mp2, bp2:
            left, right, _ = db.guide_tree(sim)
            check(lens, mp2, bp2, left, right)


def test_rejects_bad_tree():
    seqs = ou.stemloop_family(3, 30, 2)
    lens, mp, bp, sim = inputs(seqs)
    with pytest.raises(db.DafsCudaError):
        db.progressive(lens, mp, bp, [-1, -1, -1, 0, 0], [-1, -1, -1, 1, 3])     # node 0 used twice
    with pytest.raises(db.DafsCudaError):
        db.progressive(lens, mp, bp, [-1, -1, -1, 0, 4], [-1, -1, -1, 1, 2])     # child not earlier


def test_device_resident_pipeline_equals_host_pipeline():
    """Posteriors kept in HBM between the stages (fetch_device -> dafs_cuda_pct_device ->
    dafs_cuda_progressive) give the bits of the pipeline that passes host batches."""
    import dafs_b200 as db
    seqs = db.homolog_family(14, 110, config_id=31)
    lens = np.array([len(q) for q in seqs], np.int32)
    pi, pj = db.all_pairs(len(seqs))
    # host route
    bp_h = db.mccaskill_all(seqs, 0.01); mp_h = db.pairhmm_all(seqs, 0.01)
    sim_h, mp2_h, bp2_h = db.pct(lens, mp_h, bp_h, 0.25, 0.25)
    left, right, _ = db.guide_tree(sim_h)
    g_h = db.progressive(lens, mp2_h, bp2_h, left, right)
    # device route
    with db.McCaskillPlan(seqs, 0.01) as pm, db.PairHmmPlan(seqs, pi, pj, 0.01) as pp:
        pm.run(); pp.run()
        bp_d = pm.fetch_device(); mp_d = pp.fetch_device()
        with pytest.raises(db.DafsCudaError):
            pp.run()                                   # the plan gave its result arrays away
    assert mp_d.nnz == mp_h.nnz and np.array_equal(mp_d.nz_off, mp_h.nz_off)
    sim_d, mp2_d, bp2_d = db.pct(lens, mp_d, bp_d, 0.25, 0.25, device_out=True)
    assert np.array_equal(sim_d.view(np.uint32), sim_h.view(np.uint32))
    for dev, host in ((mp2_d, mp2_h), (bp2_d, bp2_h), (mp_d, mp_h)):
        with dev.to_host() as c:
            assert np.array_equal(c.row_ptr, host.row_ptr) and np.array_equal(c.col, host.col)
            assert np.array_equal(c.val.view(np.uint32), host.val.view(np.uint32))
    g_d = db.progressive(lens, mp2_d, bp2_d, left, right)
    assert g_d["L"] == g_h["L"] and np.array_equal(g_d["bits"], g_h["bits"]) and np.array_equal(g_d["ss"], g_h["ss"])
    assert g_d["score"].view(np.uint32) == g_h["score"].view(np.uint32)
    # zero weights: the outputs are copies of the (device) inputs
    _, c_mp, c_bp = db.pct(lens, mp_d, bp_d, 0.0, 0.0, device_out=True)
    with c_mp.to_host() as c:
        assert np.array_equal(c.val.view(np.uint32), mp_h.val.view(np.uint32))
    for b in (bp_d, mp_d, mp2_d, bp2_d, c_mp, c_bp, bp_h, mp_h, mp2_h, bp2_h):
        b.close()
