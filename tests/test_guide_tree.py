"""dafs_cuda_guide_tree (DAFS::build_tree, src/dafs.cpp:447-495; host arithmetic inside
libdafs_cuda, no device needed) against the restatement and the reference's own function."""
import numpy as np
import pytest

import oracle_util as ou
import dafs_b200 as db


def sims(n, seed, ties=False):
    rs = np.random.RandomState(seed)
    s = rs.rand(n, n).astype(np.float32)
    if ties:
        s = np.round(s * 4) / 4
    s = np.maximum(s, s.T)
    np.fill_diagonal(s, 1.0)
    return s.astype(np.float32)


@pytest.mark.parametrize("n,seed,ties", [(1, 0, False), (2, 1, False), (3, 2, False), (9, 3, False), (17, 4, True), (64, 5, False), (40, 6, True)])
def test_tree_matches_restatement(n, seed, ties):
    s = sims(n, seed, ties)
    left, right, score = db.guide_tree(s)
    ol, orr, osc = ou.oracle_build_tree(s)
    assert np.array_equal(left, ol) and np.array_equal(right, orr)
    assert np.array_equal(score.view(np.uint32), osc.view(np.uint32))
    assert (left[:n] == -1).all() and (left[n:] >= 0).all()
    assert sorted(np.concatenate([left[n:], right[n:]]).tolist()) == list(range(2 * n - 2))


def test_rejects_empty():
    with pytest.raises(db.DafsCudaError):
        db.lib()  # make sure the library is loaded
        from dafs_b200.capi import _check
        _check(db.lib().dafs_cuda_guide_tree(0, None, None, None, None))
