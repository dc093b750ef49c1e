"""CPU tests: the C restatement (oracle/pairhmm_oracle.c) against the golden
vectors generated from the reference's own code, against README.md:59 of the
reference, and -- when oracle/_ref is built -- bit-for-bit against it."""
import json
import os
import zlib

import numpy as np
import pytest

import oracle_util as ou
from dafs_b200.synth import random_rna

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "pairhmm_golden.json")))["cases"]


@pytest.mark.parametrize("case", GOLD, ids=[c["name"] for c in GOLD])
def test_oracle_matches_reference_golden(case):
    rp, col, val, _ = ou.oracle_pair_csr(case["s1"].encode(), case["s2"].encode(), case["th"])
    assert len(col) == case["nnz"]
    assert zlib.crc32(rp.tobytes()) == case["crc_row_ptr"]
    assert zlib.crc32(col.tobytes()) == case["crc_col"]
    assert zlib.crc32(val.tobytes()) == case["crc_val"]      # bit-exact FP32


def test_readme_guide_tree_leaf_scores():
    """README.md:59 of the reference: 0.96716, 0.935672, 0.933469, 0.74278."""
    want = {"RF00005:0,5": "0.96716", "RF00005:1,9": "0.935672",
            "RF00005:6,7": "0.933469", "RF00005:2,3": "0.74278"}
    by = {c["name"]: c for c in GOLD}
    for k, v in want.items():
        assert "%g" % by[k]["sim"] == v


@pytest.mark.skipif(not ou.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("l1,l2,seed", [(1, 1, 0), (2, 3, 1), (64, 64, 2), (97, 131, 3), (300, 300, 4), (31, 420, 5)])
def test_oracle_bit_exact_vs_reference_library(l1, l2, seed):
    s = random_rna(2, max(l1, l2), config_id=50 + seed)
    a, b = s[0][:l1], s[1][:l2]
    d_ref, _ = ou.ref_pair_dense(a, b, 0.01)
    d_orc = ou.oracle_pair_dense(a, b, 0.01)
    assert np.array_equal(d_ref.view(np.uint32), d_orc.view(np.uint32))
