"""GPU test of the allocation-free entry point dafs_cuda_nw_decode_static (include/dafs_cuda.h):
its results must equal dafs_cuda_nw_decode's bit for bit -- the tolerance of every NW test in this
repo -- on calls in a row that share the library's static device scratch.  The file name sorts
last on purpose: a failure here must not hide other tests."""
import numpy as np
import pytest

import oracle_util as ou

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dafs():
    import dafs_b200
    dafs_b200.init(1)
    return dafs_b200


def bits(x):
    return int(np.float32(x).view(np.uint32))


def problem(L1, L2, seed):
    _, _, p = ou.dd_problem(max(L1, L2), family=seed)
    p = np.ascontiguousarray(p[:L1, :L2])
    r = np.random.default_rng(seed)
    q = ((r.random((L1, L2)) < 0.05) * r.random((L1, L2)) * 0.5).astype(np.float32)
    return p, q


def test_static_scratch_decode_equals_allocating_decode_twice_in_a_row(dafs):
    # a large problem first, then smaller ones: the second and third call find the first call's
    # bytes in the scratch (stale traceback planes, a stale score and bad-cell flag)
    for L1, L2, seed in ((300, 280, 3), (73, 80, 4), (150, 301, 5), (1, 1, 0)):
        p, q = problem(L1, L2, seed)
        for qq in (q, None):
            s_ref, al_ref = dafs.nw_decode(p, qq, 0.01)
            s1, al1 = dafs.nw_decode_static(p, qq, 0.01)
            s2, al2 = dafs.nw_decode_static(p, qq, 0.01)
            assert bits(s1) == bits(s_ref) and np.array_equal(al1, al_ref)
            assert bits(s2) == bits(s_ref) and np.array_equal(al2, al_ref)
            so, alo, _, bad = ou.oracle_nw(p, qq, 0.01)
            assert bad == 0 and bits(so) == bits(s1) and np.array_equal(alo, al1)


def test_static_scratch_sparse_envelope_and_wide_band(dafs):
    r = np.random.default_rng(11)
    p = ((r.random((40, 55)) < 0.04) * r.random((40, 55))).astype(np.float32)     # rows without any cell
    s_ref, al_ref = dafs.nw_decode(p, None, 0.3)
    s1, al1 = dafs.nw_decode_static(p, None, 0.3)
    assert bits(s1) == bits(s_ref) and np.array_equal(al1, al_ref)
    p = (0.011 + 0.5 * r.random((200, 260))).astype(np.float32)                  # full-width bands (9 words)
    s_ref, al_ref = dafs.nw_decode(p, None, 0.01)
    s1, al1 = dafs.nw_decode_static(p, None, 0.01)
    assert bits(s1) == bits(s_ref) and np.array_equal(al1, al_ref)


def test_static_scratch_rejects_what_does_not_fit(dafs):
    p = np.zeros((2049, 4), np.float32)
    with pytest.raises(dafs.DafsCudaError) as e:
        dafs.nw_decode_static(p, None, 0.01)
    assert e.value.code == -1 and "static limit" in str(e.value)
    p = np.full((1500, 1500), 0.5, np.float32)          # 2.25 M envelope cells: 18 MB > 8 MiB
    with pytest.raises(dafs.DafsCudaError) as e:
        dafs.nw_decode_static(p, None, 0.01)
    assert e.value.code == -1 and "static scratch" in str(e.value)
    # and the scratch still works afterwards
    p, q = problem(64, 64, 9)
    s_ref, al_ref = dafs.nw_decode(p, q, 0.01)
    s1, al1 = dafs.nw_decode_static(p, q, 0.01)
    assert bits(s1) == bits(s_ref) and np.array_equal(al1, al_ref)
