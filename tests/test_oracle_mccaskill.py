"""CPU tests of oracle/mccaskill_oracle.c.  PARITY UNPINNED (ViennaRNA is absent,
see the oracle's header): the DP restatement is checked against exhaustive
structure enumeration under the same loop-energy functions, against
probability identities, and against small committed vectors of itself (so that
the GPU box and later rounds see the same numbers)."""
import json
import math
import os

import numpy as np
import pytest

import oracle_util as ou
from dafs_b200.synth import random_rna

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

DESIGNED = [b"GGCGAAAGCCGCGAAAGCGCC", b"GCGGCAAAAGCCGCUUUUGCGC", b"GGGCGAAAGCGCGAAAGCGCCC",
            b"GCCGGAAACCGGAAACCGGAAAGC", b"GGGGACGGGGACGGGGACGG", b"GGGAAAUCCC", b"GGUGACCC", b"ACGUN", b"A"]


@pytest.mark.parametrize("seq", DESIGNED + [random_rna(1, n, config_id=300 + n)[0] for n in (9, 13, 17, 20, 22)])
def test_dp_equals_exhaustive_enumeration(seq):
    Pb, Z = ou.oracle_mccaskill_bruteforce(seq)
    Pd, lnz = ou.oracle_mccaskill(seq)
    assert abs(math.log(Z) - lnz) < 1e-9
    assert np.abs(Pb - Pd).max() < 1e-12


@pytest.mark.parametrize("n", [40, 150, 300])
def test_probability_identities(n):
    s = random_rna(1, n, config_id=4)[0]
    P, lnz = ou.oracle_mccaskill(s)
    assert (P >= 0).all() and np.isfinite(lnz)
    assert (np.tril(P) == 0).all()
    paired = P.sum(0) + P.sum(1)
    assert paired.max() <= 1.0 + 1e-9
    for i in range(n):                       # no pair closer than TURN, none between non-complementary bases
        assert (P[i, i + 1:i + 4] == 0).all()


def test_case_and_t_equal_u():
    s = random_rna(1, 60, config_id=7)[0]
    P0, _ = ou.oracle_mccaskill(s)
    P1, _ = ou.oracle_mccaskill(s.lower().replace(b"u", b"t"))
    # tetraloop bonuses are matched on the letters, so only sequences without a
    # listed tetraloop are guaranteed equal; this one has none
    assert np.abs(P0 - P1).max() < 1e-12


def test_self_golden():
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "mccaskill_selfgolden.json")))
    for c in gold["cases"]:
        P, lnz = ou.oracle_mccaskill(c["seq"].encode())
        assert abs(lnz - c["lnz"]) < 1e-7 * max(1.0, abs(c["lnz"]))
        for i, j, p in c["top"]:
            assert abs(P[i, j] - p) < 1e-9


CONSTRAINED = [
    (b"GGCGAAAGCCGCGAAAGCGCC", b"?????????????????????"),      # no constraint at all
    (b"GGCGAAAGCCGCGAAAGCGCC", b"((?????))????????????"),      # two given pairs
    (b"GGCGAAAGCCGCGAAAGCGCC", b"..???????????????????"),      # two bases kept unpaired
    (b"GGCGAAAGCCGCGAAAGCGCC", b"(???????????????????)"),      # one pair around everything
    (b"GCGGCAAAAGCCGCUUUUGCGC", b"??(????.?)??(?..??)???"),    # pairs and unpaired bases
    (b"GGGGACGGGGACGGGGACGG", b"?(??????????????)???"),        # a pair whose ends cannot pair (G-G): it still blocks crossings
    (b"GGGAAAUCCC", b"(((????)))"),
]


@pytest.mark.parametrize("seq,con", CONSTRAINED)
def test_constrained_dp_equals_constrained_enumeration(seq, con):
    Pb, Z = ou.oracle_mccaskill_bruteforce_constrained(seq, con)
    Pd, lnz = ou.oracle_mccaskill_constrained(seq, con)
    assert abs(math.log(Z) - lnz) < 1e-9
    assert np.abs(Pb - Pd).max() < 1e-12
    n = len(seq)
    for i in range(n):
        if con[i:i + 1] == b".":
            assert Pd[i, :].max() == 0 and Pd[:, i].max() == 0


def test_free_constraint_is_the_unconstrained_ensemble():
    s = random_rna(1, 80, config_id=21)[0]
    P0, z0 = ou.oracle_mccaskill(s)
    P1, z1 = ou.oracle_mccaskill_constrained(s, b"?" * len(s))
    assert z0 == z1 and np.array_equal(P0, P1)
