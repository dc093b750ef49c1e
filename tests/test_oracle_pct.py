"""CPU: the C restatement of the consistency transformation (oracle/pct_oracle.c)
against the reference's own transpose_mp / calculate_similarity_score /
relax_basepairing_probability / relax_matching_probability compiled from
/root/reference (oracle/_ref), bit for bit; and against the committed golden."""
import json
import os

import numpy as np
import pytest

import oracle_util as ou
from dafs_b200 import homolog_family, random_rna

HERE = os.path.dirname(os.path.abspath(__file__))
KEYS = ("sim", "mp_row_ptr", "mp_col", "mp_val", "bp_row_ptr", "bp_col", "bp_val")


def same(a, b):
    for k in KEYS:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape, k
        if x.dtype == np.float32:
            assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), k
        else:
            assert np.array_equal(x, y), k


CASES = [
    ("family", 5, 40, 0.25, 0.25),
    ("family", 4, 60, 0.25, 0.25),
    ("random", 4, 35, 0.25, 0.25),
    ("family", 3, 30, 0.25, 0.25),       # N - 2 = 1
    ("family", 4, 30, -1.0, -1.0),       # negative weights: the 1/N branch
    ("family", 4, 30, 0.5, 0.0),         # base-pair relaxation off
    ("family", 4, 30, 0.0, 0.7),         # matching relaxation off
    ("ragged", 5, 30, 0.25, 0.25),
]


def make(kind, n, L, seed=11):
    if kind == "family":
        return homolog_family(n, L, config_id=seed)
    if kind == "ragged":
        s = homolog_family(n, L + 12, config_id=seed)
        return [x[: L - 5 + 4 * k] for k, x in enumerate(s)]
    return random_rna(n, L, config_id=seed)


@pytest.mark.skipif(not ou.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("kind,n,L,wa,ws", CASES)
def test_oracle_matches_reference(kind, n, L, wa, ws):
    lens, mp, bp = ou.pct_inputs(make(kind, n, L))
    same(ou.oracle_pct(lens, mp, bp, wa, ws), ou.ref_pct(lens, mp, bp, wa, ws))


def test_oracle_matches_golden():
    g = json.load(open(os.path.join(HERE, "golden", "pct_golden.json")))
    for case in g["cases"]:
        lens, mp, bp = ou.pct_inputs([s.encode() for s in case["seqs"]])
        o = ou.oracle_pct(lens, mp, bp, case["w_pct_a"], case["w_pct_s"])
        for k in KEYS:
            x = np.asarray(o[k])
            want = np.asarray(case[k], dtype=np.uint32 if x.dtype == np.float32 else x.dtype)
            got = x.view(np.uint32) if x.dtype == np.float32 else x
            assert np.array_equal(got.ravel(), want.ravel()), (case["name"], k)


def test_relaxed_matrices_are_well_formed():
    lens, mp, bp = ou.pct_inputs(make("family", 5, 40))
    o = ou.oracle_pct(lens, mp, bp)
    assert np.all(o["mp_val"] > 0.01) and np.all(o["bp_val"] > 0.01)
    assert np.allclose(np.diag(o["sim"]), 1.0)
    assert np.array_equal(o["sim"], o["sim"].T)
