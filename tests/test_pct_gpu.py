"""GPU: dafs_cuda_pct (transpose_mp, calculate_similarity_score,
relax_basepairing_probability, relax_matching_probability; src/dafs.cpp:1796-1827)
through the C ABI, bit for bit against oracle/pct_oracle.c (itself pinned against the
reference's own functions, tests/test_oracle_pct.py) and the committed golden."""
import json
import os

import numpy as np
import pytest

import oracle_util as ou
import dafs_b200 as db
from dafs_b200 import homolog_family, random_rna

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def bits(x):
    return np.asarray(x, np.float32).view(np.uint32)


def check(lens, mp, bp, wa, ws, want=None):
    o = want if want is not None else ou.oracle_pct(lens, mp, bp, wa, ws)
    sim, gmp, gbp = db.pct(lens, mp, bp, wa, ws)
    with gmp, gbp:
        assert np.array_equal(bits(sim), bits(o["sim"])), "sim"
        for name, g in (("mp", gmp), ("bp", gbp)):
            assert np.array_equal(np.asarray(g.row_ptr), np.asarray(o[name + "_row_ptr"], np.int32)), name + " row_ptr"
            assert np.array_equal(np.asarray(g.col), np.asarray(o[name + "_col"], np.uint32)), name + " col"
            assert np.array_equal(bits(g.val), bits(o[name + "_val"])), name + " val"
            assert int(g.nz_off[-1]) == g.nnz == len(o[name + "_col"])
            # nz_off agrees with the row_ptr blocks
            for u in range(g.n_units):
                rp, col, _ = g.unit(u)
                assert rp[0] == 0 and rp[-1] == len(col)


CASES = [
    ("family", 5, 40, 0.25, 0.25),
    ("family", 4, 60, 0.25, 0.25),
    ("random", 4, 35, 0.25, 0.25),
    ("family", 3, 30, 0.25, 0.25),
    ("family", 2, 30, 0.25, 0.25),       # N - 2 = 0: only z in {x, y}
    ("family", 4, 30, -1.0, -1.0),
    ("family", 4, 30, 0.5, 0.0),
    ("family", 4, 30, 0.0, 0.7),
    ("family", 4, 30, 0.0, 0.0),         # transposes and similarity only
    ("ragged", 5, 30, 0.25, 0.25),
    ("family", 12, 70, 0.25, 0.25),
    ("family", 6, 130, 0.25, 0.25),      # rows longer than one warp pass, several 32-column chunks
]


def make(kind, n, L, seed=11):
    if kind == "family":
        return homolog_family(n, L, config_id=seed)
    if kind == "ragged":
        s = homolog_family(n, L + 12, config_id=seed)
        return [x[: L - 5 + 4 * k] for k, x in enumerate(s)]
    return random_rna(n, L, config_id=seed)


@pytest.mark.parametrize("kernel", ["sparse", "dense"])
@pytest.mark.parametrize("kind,n,L,wa,ws", CASES)
def test_pct_matches_oracle(kind, n, L, wa, ws, kernel, monkeypatch):
    """Both relax_matching_probability kernels (thread per row through the row heads /
    eight lanes per row with shared-memory accumulators) on every case."""
    monkeypatch.setenv("DAFS_PCT_MP_KERNEL", kernel)
    lens, mp, bp = ou.pct_inputs(make(kind, n, L))
    check(lens, mp, bp, wa, ws)


def test_pct_matches_golden():
    g = json.load(open(os.path.join(HERE, "golden", "pct_golden.json")))
    for case in g["cases"]:
        lens, mp, bp = ou.pct_inputs([s.encode() for s in case["seqs"]])
        want = {k: np.asarray(case[k], np.uint32).view(np.float32) if k.endswith("val") or k == "sim"
                else np.asarray(case[k]) for k in ("sim", "mp_row_ptr", "mp_col", "mp_val", "bp_row_ptr", "bp_col", "bp_val")}
        want["sim"] = want["sim"].reshape(len(lens), len(lens))
        check(lens, mp, bp, case["w_pct_a"], case["w_pct_s"], want)


def test_pct_single_sequence():
    lens, mp, bp = ou.pct_inputs(make("family", 1, 30))
    check(lens, mp, bp, 0.25, 0.25)


@pytest.mark.parametrize("kernel", ["sparse", "dense", "dg", "auto"])
def test_pct_dense_rows(kernel, monkeypatch):
    """A low threshold gives rows with many entries: several k per row (more than the
    two of a row head, more than the eight of one chunk), rows of mp_[z][y] with more
    than eight entries, equal columns reached through different k."""
    if kernel != "auto":
        monkeypatch.setenv("DAFS_PCT_MP_KERNEL", kernel)    # "dg": dense kernel, accumulators in global scratch
    seqs = make("random", 4, 48)
    lens, mp, bp = ou.pct_inputs(seqs, th_a=1e-4, th_bp=1e-4)
    assert mp.nz_off[-1] > 9 * 48 * 6
    check(lens, mp, bp, 0.25, 0.25)


def test_pct_from_device_results():
    """The hand-over a caller makes: pair-HMM and McCaskill batches straight into pct."""
    seqs = make("family", 6, 50)
    lens = np.array([len(s) for s in seqs], np.int32)
    with db.pairhmm_all(seqs, 0.01) as mp, db.mccaskill_all(seqs, 0.01) as bp:
        want = ou.oracle_pct(lens, mp, bp, 0.25, 0.25)
        check(lens, mp, bp, 0.25, 0.25, want)


def test_pct_rejects_malformed_input():
    lens, mp, bp = ou.pct_inputs(make("family", 3, 30))
    bad = ou.FlatCsr.__new__(ou.FlatCsr)
    bad.row_off, bad.nz_off, bad.row_ptr = mp.row_off, mp.nz_off, mp.row_ptr
    bad.val = mp.val
    bad.col = mp.col.copy(); bad.col[0] = 10_000
    with pytest.raises(db.DafsCudaError):
        db.pct(lens, bad, bp)
    with pytest.raises(db.DafsCudaError):
        db.pct(lens[:2], mp, bp)
