"""GPU parity tests of path (1): libdafs_cuda through its C ABI against the
golden vectors from the reference and against the C oracle on seeded inputs.
Bar: bit-exact CSR (the kernels reproduce the reference's unfused FP32
operation order), which is stricter than north_star's 1e-4 absolute."""
import json
import os
import zlib

import numpy as np
import pytest

import dafs_b200
import oracle_util as ou
from dafs_b200.synth import fasta_read, homolog_family, random_rna

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "pairhmm_golden.json")))["cases"]


def check_batch_against_oracle(seqs, pi, pj, th, batch):
    assert batch.n_units == len(pi)
    for u, (i, j) in enumerate(zip(pi, pj)):
        rp, col, val = batch.unit(u)
        orp, ocol, oval, _ = ou.oracle_pair_csr(seqs[i], seqs[j], th)
        assert np.array_equal(rp, orp), "row_ptr differs for pair %d (%d,%d)" % (u, i, j)
        assert np.array_equal(col, ocol), "col differs for pair %d" % u
        assert np.array_equal(val.view(np.uint32), oval.view(np.uint32)), \
            "val differs for pair %d: max abs %g" % (u, np.abs(val - oval).max())


def test_golden_cases_bit_exact():
    """Every golden case (reference-generated) as a batch of one."""
    for c in GOLD:
        with dafs_b200.pairhmm_pairs([c["s1"].encode(), c["s2"].encode()], [0], [1], c["th"]) as b:
            rp, col, val = b.unit(0)
            assert len(col) == c["nnz"], c["name"]
            assert zlib.crc32(rp.tobytes()) == c["crc_row_ptr"], c["name"]
            assert zlib.crc32(col.tobytes()) == c["crc_col"], c["name"]
            assert zlib.crc32(val.tobytes()) == c["crc_val"], c["name"]


def test_rf00005_all_pairs_one_call(golden_dir):
    _, seqs = fasta_read(os.path.join(golden_dir, "RF00005_0.fa"))
    by = {c["name"]: c for c in GOLD}
    with dafs_b200.pairhmm_all(seqs, 0.01) as b:
        pi, pj = dafs_b200.all_pairs(len(seqs))
        assert b.n_units == 45
        for u, (i, j) in enumerate(zip(pi, pj)):
            c = by["RF00005:%d,%d" % (i, j)]
            rp, col, val = b.unit(u)
            assert zlib.crc32(col.tobytes()) == c["crc_col"]
            assert zlib.crc32(val.tobytes()) == c["crc_val"]


@pytest.mark.parametrize("lens,th", [
    ([1, 2, 3, 5, 8, 13], 0.01),                 # tiny / ragged, C=2
    ([60, 63, 64, 65, 95, 96, 97], 0.01),        # stripe-width edges for small C
    ([150, 149, 151, 120], 0.01),                # config C3 length
    ([300, 300, 299, 301], 0.01),                # config C4 length, C=10
    ([319, 320, 321, 383, 384, 385], 0.01),      # one / two stripes
    ([40, 700, 1100], 0.01),                     # long columns: 3 stripes of C=12
    ([150, 160], 0.2), ([150, 160], 0.0004), ([90, 80], 0.0),
])
def test_seeded_batches_match_oracle(lens, th):
    pool = random_rna(len(lens), max(lens), config_id=7)
    seqs = [s[:l] for s, l in zip(pool, lens)]
    pi, pj = dafs_b200.all_pairs(len(seqs))
    # both orientations: rows/columns swap roles
    pi2 = np.concatenate([pi, pj]); pj2 = np.concatenate([pj, pi])
    with dafs_b200.pairhmm_pairs(seqs, pi2, pj2, th) as b:
        check_batch_against_oracle(seqs, pi2, pj2, th, b)


def test_homologs_and_alphabet():
    fam = homolog_family(5, 200, config_id=8)
    odd = [b"acgutnACGUTNxX-acguACGU", b"ACGTTNnnacguRYKMacgtACGU", b"NNNNNNNN", b"T", b"u"]
    seqs = fam + odd
    pi, pj = dafs_b200.all_pairs(len(seqs))
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
        check_batch_against_oracle(seqs, pi, pj, 0.01, b)


def test_plan_rerun_is_idempotent_and_timed():
    seqs = random_rna(24, 150, config_id=3)
    pi, pj = dafs_b200.all_pairs(len(seqs))
    with dafs_b200.PairHmmPlan(seqs, pi, pj, 0.01) as plan:
        plan.run(); ms1, nnz1 = plan.sync()
        with plan.fetch() as b1:
            v1 = b1.val.copy(); c1 = b1.col.copy(); r1 = b1.row_ptr.copy()
        plan.run(); ms2, nnz2 = plan.sync()
        with plan.fetch() as b2:
            assert nnz1 == nnz2 == b2.nnz
            assert np.array_equal(v1.view(np.uint32), b2.val.view(np.uint32))
            assert np.array_equal(c1, b2.col) and np.array_equal(r1, b2.row_ptr)
        assert ms1 > 0 and ms2 > 0


def test_pool_overflow_path_regrows():
    """A tiny threshold keeps far more entries than the density guess."""
    seqs = random_rna(6, 120, config_id=9)
    pi, pj = dafs_b200.all_pairs(len(seqs))
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 1e-6) as b:
        assert b.nnz > 10 * 120 * len(pi)
        check_batch_against_oracle(seqs, pi, pj, 1e-6, b)


def test_full_size_properties_c4_sample():
    """Config C4 lengths at a batch the oracle cannot check exhaustively:
    size-independent properties + a sampled exact check."""
    seqs = random_rna(48, 300, config_id=4)
    pi, pj = dafs_b200.all_pairs(len(seqs))
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
        assert b.n_units == 1128
        assert b.val.min() > 0.01 and b.val.max() <= 1.0
        # ascending columns inside every row; posterior rows sum to <= 1
        for u in range(0, b.n_units, 37):
            rp, col, val = b.unit(u)
            for r in range(300):
                c = col[rp[r]:rp[r + 1]]
                assert np.all(np.diff(c.astype(np.int64)) > 0)
                assert val[rp[r]:rp[r + 1]].sum() <= 1.01   # the reference's EXP/LOOKUP pieces overshoot a little
        sample = list(range(0, b.n_units, 97))
        check = dafs_b200.pairhmm_pairs(seqs, pi[sample], pj[sample], 0.01)
        for k, u in enumerate(sample):
            a = b.unit(u); c = check.unit(k)
            assert all(np.array_equal(x, y) for x, y in zip(a, c))
            orp, ocol, oval, _ = ou.oracle_pair_csr(seqs[pi[u]], seqs[pj[u]], 0.01)
            assert np.array_equal(a[1], ocol) and np.array_equal(a[2].view(np.uint32), oval.view(np.uint32))
        check.close()


def test_bad_arguments():
    with pytest.raises(dafs_b200.DafsCudaError):
        dafs_b200.pairhmm_pairs([b"ACGU", b""], [0], [1], 0.01)      # empty sequence
    with pytest.raises(dafs_b200.DafsCudaError):
        dafs_b200.pairhmm_pairs([b"ACGU", b"AC"], [0], [2], 0.01)    # pair index out of range
    with dafs_b200.pairhmm_pairs([b"ACGU"], [], [], 0.01) as b:      # empty batch
        assert b.n_units == 0 and b.nnz == 0


def _same_batch(a, b):
    assert a.nnz == b.nnz
    assert np.array_equal(np.asarray(a.row_off), np.asarray(b.row_off))
    assert np.array_equal(np.asarray(a.nz_off), np.asarray(b.nz_off))
    assert np.array_equal(np.asarray(a.row_ptr), np.asarray(b.row_ptr))
    assert np.array_equal(np.asarray(a.col), np.asarray(b.col))
    assert np.array_equal(np.asarray(a.val).view(np.uint32), np.asarray(b.val).view(np.uint32))


@pytest.mark.parametrize("th", [0.01, 1e-5])
def test_chunked_fetch_equals_single_launch(th):
    """The one-call entry point runs the pair list as several launches and copies finished
    chunks to the host under the next launch; the split-phase plan runs one launch.  Same
    bits either way (th = 1e-5 overflows the first pool estimate: regrow + rerun path), and
    equal to the oracle on a sample of pairs."""
    import dafs_b200 as db
    seqs = db.random_rna(180, 30, config_id=31)
    pi, pj = db.all_pairs(len(seqs))
    assert len(pi) >= 6 * 148 * 16                      # enough rounds of resident warps to be chunked
    plan = db.PairHmmPlan(seqs, pi, pj, th)
    plan.run()
    with plan.fetch() as single, db.pairhmm_pairs(seqs, pi, pj, th) as chunked:
        assert plan.info()["launches_per_run"] == 3
        _same_batch(single, chunked)
        for u in range(0, len(pi), 331):
            rp, col, val = chunked.unit(u)
            orp, ocol, oval, _ = ou.oracle_pair_csr(seqs[pi[u]], seqs[pj[u]], th)
            assert np.array_equal(rp, orp) and np.array_equal(col, ocol)
            assert np.array_equal(val.view(np.uint32), oval.view(np.uint32))
    plan.close()


def test_config_c5_length_pairs_bit_exact():
    """Config C5 (L = 1500, 2.25 M cells per pair, 5 stripes): every pair of a 1500 / 1500 /
    1499 triple in both orientations, plus a 40 x 1500 strip and its transpose."""
    pool = random_rna(4, 1500, config_id=5)
    seqs = [pool[0], pool[1], pool[2][:1499], pool[3][:40]]
    pi = np.array([0, 0, 1, 1, 2, 2, 3, 0], np.int32)
    pj = np.array([1, 2, 2, 0, 0, 1, 0, 3], np.int32)
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
        check_batch_against_oracle(seqs, pi, pj, 0.01, b)
        assert b.nnz > 6 * 1499


def test_config_c5_homologs_bit_exact():
    fam = homolog_family(3, 1500, config_id=5)
    pi, pj = dafs_b200.all_pairs(3)
    with dafs_b200.pairhmm_pairs(fam, pi, pj, 0.01) as b:
        check_batch_against_oracle(fam, pi, pj, 0.01, b)


def test_rows_beyond_shared_counters_bit_exact():
    """More than 2048 rows: the per-row counters of the CSR emission leave shared memory for the forward plane
    (DAFS_PAIRHMM_CNT_ROWS); 2100 x 300 and 300 x 2100 also put eight chained stripes behind two short ones."""
    fam = homolog_family(2, 2100, config_id=7)
    seqs = [fam[0], fam[1][:2047], fam[1][:300]]
    pi = np.array([0, 1, 0, 2, 1], np.int32)
    pj = np.array([1, 0, 2, 0, 2], np.int32)
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
        check_batch_against_oracle(seqs, pi, pj, 0.01, b)


@pytest.mark.parametrize("cols", [2, 3, 4, 5, 6, 8, 10, 12])
def test_every_stripe_width_bit_exact(cols, monkeypatch):
    """Each compiled columns-per-lane variant (stripes of 64..384 columns), lengths around the stripe width and
    around the minimum period of the chained stripes (short rows are spaced by idle steps)."""
    monkeypatch.setenv("DAFS_PAIRHMM_COLS", str(cols))
    sw = 32 * cols
    pool = random_rna(6, 2 * sw + 40, config_id=cols)
    seqs = [pool[0][:sw - 1], pool[1][:sw], pool[2][:sw + 1], pool[3], pool[4][:97], pool[5][:1]]
    pi, pj = dafs_b200.all_pairs(len(seqs))
    pi = np.concatenate([pi, pj]).astype(np.int32)
    pj = np.concatenate([pj, pi[:len(pj)]]).astype(np.int32)
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
        check_batch_against_oracle(seqs, pi, pj, 0.01, b)
