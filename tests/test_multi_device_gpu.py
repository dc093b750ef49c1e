"""GPU tests of the in-process multi-device route of the C ABI (ngpu > 1: one host thread per
device, host merge -- what the single-process `dafs` binary calls): its result must equal the
one-device result bit for bit.  Skipped when fewer than two devices are visible."""
import numpy as np
import pytest

from dafs_b200.synth import random_rna

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dafs():
    import dafs_b200
    if dafs_b200.device_count() < 2:
        pytest.skip("needs two visible devices")
    dafs_b200.init(0)
    return dafs_b200


def same(a, b):
    assert a.n_units == b.n_units and a.nnz == b.nnz
    assert np.array_equal(a.row_off, b.row_off) and np.array_equal(a.nz_off, b.nz_off)
    assert np.array_equal(a.row_ptr, b.row_ptr) and np.array_equal(a.col, b.col)
    assert np.array_equal(a.val.view(np.uint32), b.val.view(np.uint32))


@pytest.mark.parametrize("ngpu", [2, 0])
def test_pairhmm_all_two_devices_equals_one(dafs, ngpu):
    seqs = [s[:110 + 7 * k] for k, s in enumerate(random_rna(24, 300, config_id=2))]
    n = dafs.device_count() if ngpu == 0 else ngpu
    with dafs.pairhmm_all(seqs, 0.01, 1) as one, dafs.pairhmm_all(seqs, 0.01, n) as many:
        assert one.n_units == 276
        same(one, many)


def test_pairhmm_all_repeated_threaded_calls(dafs):
    """The per-device lazy set-up (constant tables, kernel attributes) is reached from two host
    threads at once on the first call; later calls reuse it."""
    seqs = random_rna(12, 90, config_id=12)
    with dafs.pairhmm_all(seqs, 0.01, 1) as one:
        for _ in range(3):
            with dafs.pairhmm_all(seqs, 0.01, 2) as two:
                same(one, two)


def test_mccaskill_all_two_devices_equals_one(dafs):
    seqs = [s[:60 + 11 * k] for k, s in enumerate(random_rna(13, 220, config_id=3))]
    with dafs.mccaskill_all(seqs, 0.01, 1) as one, dafs.mccaskill_all(seqs, 0.01, 2) as two:
        assert one.n_units == 13
        same(one, two)
