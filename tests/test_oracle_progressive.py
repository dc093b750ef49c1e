"""The C restatement of the progressive stage (oracle/prog_oracle.c) against the reference's
own build_tree / average_* / project_* / align_alignments / align (lifted from
/root/reference/src/dafs.cpp into oracle/_ref at build time), bit for bit, and against
the committed golden fixture (made by tests/golden/make_prog_golden.py)."""
import json
import os

import numpy as np
import pytest

import oracle_util as ou
from dafs_b200 import synth

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "prog_golden.json")


def family(n, length, seed, mutation=0.12):
    return ou.stemloop_family(n, length, seed, mutation)


def pipeline(seqs, w_pct=0.25):
    lens, mp, bp = ou.pct_inputs(seqs)
    r = ou.oracle_pct(lens, mp, bp, w_pct, w_pct)
    mp2, bp2 = ou.pct_to_flat(lens, r)
    return lens, mp2, bp2, r["sim"]


def same(a, b):
    assert a["L"] == b["L"]
    assert np.array_equal(a["order"], b["order"])
    assert np.array_equal(a["bits"], b["bits"])
    assert np.array_equal(a["ss"], b["ss"])
    assert np.array_equal(a["log"], b["log"])
    assert np.array_equal(a["log_score"].view(np.uint32), b["log_score"].view(np.uint32))
    assert a["score"].view(np.uint32) == b["score"].view(np.uint32)


CASES = [(2, 40, 1), (3, 50, 2), (5, 60, 3), (8, 45, 4), (6, 70, 5)]


@pytest.mark.skipif(not ou.have_ref(), reason="oracle/_ref not built (no /root/reference)")
@pytest.mark.parametrize("n,length,seed", CASES)
def test_restatement_matches_reference(n, length, seed):
    seqs = family(n, length, seed)
    lens, mp, bp, sim = pipeline(seqs)
    r = ou.ref_progressive(lens, mp, bp, sim=sim)                      # the reference builds the tree
    left, right, score = ou.oracle_build_tree(sim)
    assert np.array_equal(left, r["left"]) and np.array_equal(right, r["right"])
    assert np.array_equal(score.view(np.uint32), r["tree_score"].view(np.uint32))
    o = ou.oracle_progressive(lens, mp, bp, (left, right))
    same(o, r)


@pytest.mark.skipif(not ou.have_ref(), reason="oracle/_ref not built (no /root/reference)")
def test_caterpillar_and_unrelated():
    # a chain-shaped tree and unrelated sequences (many iterations, long alignments)
    seqs = synth.random_rna(5, 48, config_id=77)
    lens, mp, bp, sim = pipeline(seqs)
    n = len(seqs)
    left = np.full(2 * n - 1, -1, np.int32); right = np.full(2 * n - 1, -1, np.int32)
    left[n], right[n] = 3, 1
    for k in range(n + 1, 2 * n - 1):
        left[k], right[k] = [0, 2, 4][k - n - 1], k - 1                 # leaf on the left: mp_[a][b] with a > b as well
    o = ou.oracle_progressive(lens, mp, bp, (left, right), t_max=200)
    r = ou.ref_progressive(lens, mp, bp, tree=(left, right), t_max=200)
    same(o, r)


def test_single_sequence():
    seqs = synth.random_rna(1, 30, config_id=5)
    lens, mp, bp = ou.pct_inputs(seqs)
    o = ou.oracle_progressive(lens, mp, bp, (np.array([-1], np.int32), np.array([-1], np.int32)))
    assert o["L"] == 30 and o["bits"].all() and o["score"] == 0.0


def test_golden():
    g = json.load(open(GOLDEN))
    for case in g["cases"]:
        seqs = [s.encode() for s in case["seqs"]]
        lens, mp, bp, sim = pipeline(seqs)
        left, right, score = ou.oracle_build_tree(sim)
        assert left.tolist() == case["left"] and right.tolist() == case["right"]
        o = ou.oracle_progressive(lens, mp, bp, (left, right))
        assert o["L"] == case["L"]
        assert o["order"].tolist() == case["order"]
        assert ["".join("1" if b else "0" for b in row) for row in o["bits"]] == case["bits"]
        assert [int(np.int32(v)) for v in o["ss"].view(np.int32)] == case["ss"]
        assert o["log"].tolist() == case["log"]
        assert o["log_score"].view(np.uint32).tolist() == case["log_score_bits"]
        assert int(o["score"].view(np.uint32)) == case["score_bits"]
