"""GPU parity tests of (2): mccaskill_kernel against the FP64 restatement
(oracle/mccaskill_oracle.c; parity with the reference itself is UNPINNED, see its
header).  Tolerance: |p_gpu - p_oracle| <= 1e-4 absolute (BASELINE.json
north_star); observed differences are ~1e-6 (FP32-stored Boltzmann factors)."""
import os

import numpy as np
import pytest

import oracle_util as ou
from dafs_b200.synth import fasta_read, random_rna

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-4


@pytest.fixture(scope="module")
def dafs():
    import dafs_b200
    dafs_b200.init(1)
    return dafs_b200


def check_batch(dafs, seqs, th=0.01, oracle_results=None):
    with dafs.McCaskillPlan(seqs, th) as plan:
        plan.run()
        b, lnz = plan.fetch()
        with b:
            assert b.n_units == len(seqs)
            worst = 0.0
            for u, s in enumerate(seqs):
                n = len(s)
                P, lz = oracle_results[u] if oracle_results else ou.oracle_mccaskill(s)
                assert abs(lnz[u] - lz) < 1e-5 * max(1.0, abs(lz)), (u, lnz[u], lz)
                rp, col, val = b.unit(u)
                assert len(rp) == n + 1 and rp[0] == 0 and rp[-1] == len(col)
                rows = np.repeat(np.arange(n), np.diff(rp))
                assert (col > rows).all()                               # upper triangle
                for r in range(n):                                      # ascending columns
                    c = col[rp[r]:rp[r + 1]]
                    assert (np.diff(c.astype(np.int64)) > 0).all()
                assert (val > th).all()
                G = np.zeros((n, n))
                G[rows, col] = val
                keep_o = P.astype(np.float32) > th
                # every entry either side keeps must agree within TOL; entries within TOL
                # of the threshold may legitimately differ in membership
                diff = np.abs(G - np.where(keep_o, P, 0.0))
                near = np.abs(P - th) <= TOL
                assert diff[~near].max(initial=0.0) <= TOL
                worst = max(worst, np.abs(G - P)[G > 0].max(initial=0.0))
            return worst


def test_short_and_designed(dafs):
    seqs = [b"GGCGAAAGCCGCGAAAGCGCC", b"GCGGCAAAAGCCGCUUUUGCGC", b"GGGGACGGGGACGGGGACGG", b"GGGAAAUCCC", b"A", b"ACGU",
            b"ACGUNNACGU", b"gggaaauccc", b"GGGAAATCCC"]
    assert check_batch(dafs, seqs) <= TOL


def test_rfam_examples(dafs):
    for fa in ("RF00005_0.fa", "RF00017_4.fa"):
        seqs = fasta_read(os.path.join(ROOT, "tests", "golden", fa))[1]
        assert check_batch(dafs, seqs) <= TOL


@pytest.mark.parametrize("n,count", [(73, 6), (150, 6), (300, 4), (601, 2)])
def test_random_lengths(dafs, n, count):
    seqs = random_rna(count, n, config_id=3)
    assert check_batch(dafs, seqs) <= TOL


def test_ragged_batch_and_thresholds(dafs):
    r = random_rna(8, 200, config_id=9)
    seqs = [r[k][:20 + 25 * k] for k in range(8)]
    for th in (0.01, 0.2, 0.001):
        assert check_batch(dafs, seqs, th) <= TOL


def test_gc_rich_needs_rescaling(dafs):
    """ln Z / n ~ 1: the first inside pass ends far from 1 and the kernel reruns
    with the ensemble's own scale."""
    s = (b"GGGGCCCCGGGGAAAACCCCGGGGCCCC" * 20)[:500]
    assert check_batch(dafs, [s]) <= TOL


def test_all_entry_point_matches_plan(dafs):
    seqs = random_rna(5, 90, config_id=11)
    with dafs.mccaskill_all(seqs, 0.01) as a, dafs.McCaskillPlan(seqs, 0.01) as plan:
        plan.run()
        b, _ = plan.fetch()
        with b:
            assert np.array_equal(a.row_ptr, b.row_ptr) and np.array_equal(a.col, b.col)
            assert np.array_equal(a.val.view(np.uint32), b.val.view(np.uint32))


def test_config_c5_length(dafs):
    """Config C5 (L = 1500): two random sequences and a GC-rich 1500-mer whose first inside pass
    ends far from 1, so the kernel reruns with the ensemble's own scale (the rescaling path
    grows with L).  The FP64 restatement needs ~40 s per sequence: the three run side by side."""
    from concurrent.futures import ThreadPoolExecutor
    seqs = random_rna(2, 1500, config_id=5) + [(b"GGGGCCCCGGGGAAAACCCCGGGGCCCC" * 60)[:1500]]
    with ThreadPoolExecutor(3) as ex:
        oracle_results = list(ex.map(ou.oracle_mccaskill, seqs))
    assert oracle_results[2][1] / 1500 > 0.8          # ln Z per nucleotide far above the first-pass scale
    assert check_batch(dafs, seqs, 0.01, oracle_results) <= TOL


def test_repeatable(dafs):
    """Run-to-run identity of the kernel (dynamic cell hand-out, rolling rows, 8-lane groups): the
    order in which groups pick up cells varies between runs, the bits of the result must not."""
    seqs = random_rna(6, 120, config_id=13) + random_rna(3, 520, config_id=14)
    with dafs.McCaskillPlan(seqs, 0.01) as plan:
        first = None
        for rep in range(12):
            plan.run()
            b, lnz = plan.fetch()
            with b:
                cur = (b.row_ptr.copy(), b.col.copy(), b.val.view(np.uint32).copy(), np.array(lnz))
            if first is None:
                first = cur
            else:
                assert all(np.array_equal(x, y) for x, y in zip(first, cur)), rep


def constrained_case(seq, frac_pairs=0.5, seed=0):
    """A constraint in the reference's alphabet derived from the sequence's own likely structure:
    some pairs of the unconstrained ensemble become '(' ')', a few other bases '.', the rest '?'
    -- what update_basepairing_probability builds from a predicted structure (dafs.cpp:633-652)."""
    r = np.random.default_rng(seed)
    P, _ = ou.oracle_mccaskill(seq)
    n = len(seq)
    con = [b"?"] * n
    used = np.zeros(n, bool)
    order = np.dstack(np.unravel_index(np.argsort(-P, axis=None), P.shape))[0]
    chosen = []
    for i, j in order[:4 * n]:
        if P[i, j] < 0.3:
            break
        if used[i] or used[j] or any((a < i < b < j) or (i < a < j < b) for a, b in chosen):
            continue
        if r.random() < frac_pairs:
            chosen.append((i, j)); used[i] = used[j] = True
            con[i], con[j] = b"(", b")"
    for k in range(n):
        if not used[k] and r.random() < 0.1:
            con[k] = b"."
    return b"".join(con)


def check_constrained(dafs, seqs, cons, th=0.01):
    worst = 0.0
    with dafs.mccaskill_constrained(seqs, cons, th) as b:
        for u, (s, c) in enumerate(zip(seqs, cons)):
            n = len(s)
            P, _ = ou.oracle_mccaskill(s) if c is None else ou.oracle_mccaskill_constrained(s, c)
            rp, col, val = b.unit(u)
            rows = np.repeat(np.arange(n), np.diff(rp))
            G = np.zeros((n, n)); G[rows, col] = val
            keep_o = P.astype(np.float32) > th
            diff = np.abs(G - np.where(keep_o, P, 0.0))
            near = np.abs(P - th) <= TOL
            assert diff[~near].max(initial=0.0) <= TOL, (u, diff[~near].max())
            worst = max(worst, np.abs(G - P)[G > 0].max(initial=0.0))
    return worst


def test_constrained_short_cases(dafs):
    """Row f4 (src/fold.cpp:118-166): the cases of tests/test_oracle_mccaskill.py, where the restatement
    equals exhaustive enumeration under the same constraint."""
    cases = [(b"GGCGAAAGCCGCGAAAGCGCC", b"?????????????????????"), (b"GGCGAAAGCCGCGAAAGCGCC", b"((?????))????????????"),
             (b"GGCGAAAGCCGCGAAAGCGCC", b"..???????????????????"), (b"GGCGAAAGCCGCGAAAGCGCC", b"(???????????????????)"),
             (b"GCGGCAAAAGCCGCUUUUGCGC", b"??(????.?)??(?..??)???"), (b"GGGGACGGGGACGGGGACGG", b"?(??????????????)???"),
             (b"GGGAAAUCCC", b"(((????)))"), (b"GGGAAAUCCC", None)]
    seqs, cons = [c[0] for c in cases], [c[1] for c in cases]
    assert check_constrained(dafs, seqs, cons, 0.001) <= TOL


def test_constrained_rfam_and_random(dafs):
    seqs = fasta_read(os.path.join(ROOT, "tests", "golden", "RF00005_0.fa"))[1][:5] + \
        fasta_read(os.path.join(ROOT, "tests", "golden", "RF00017_4.fa"))[1][:3] + random_rna(2, 400, config_id=17)
    seqs = [q if isinstance(q, bytes) else q.encode() for q in seqs]
    cons = [constrained_case(q, seed=k) for k, q in enumerate(seqs)]
    assert sum(c.count(b"(") for c in cons) >= 20
    assert check_constrained(dafs, seqs, cons) <= TOL
    # constraints change the answer: the same batch without them differs somewhere
    with dafs.mccaskill_constrained(seqs, cons) as a, dafs.mccaskill_all(seqs) as b:
        assert a.nnz != b.nnz or not np.array_equal(a.val, b.val)


def test_constraint_rejected_when_unbalanced(dafs):
    with pytest.raises(dafs.DafsCudaError):
        dafs.mccaskill_constrained([b"GGGAAAUCCC"], [b"((??????)?"])


# ---- the banded kernel (mccaskill_kernel<GL, true>): one CTA per sequence --------------------------------
# Small batches get a thread-block cluster per sequence and the streaming kernel, so the cases above never
# reach the band pass; these do: DAFS_MCCASKILL_CLUSTER=1 gives every sequence one CTA whatever the batch, and
# the last case is the throughput form itself (more sequences than SMs: two CTAs of 384 threads per SM).

@pytest.mark.parametrize("lanes", [8, 16, 32])
@pytest.mark.parametrize("n,count", [(47, 3), (150, 3), (300, 3), (601, 2)])
def test_banded_kernel_random_lengths(dafs, monkeypatch, n, count, lanes):
    """Lengths below, at and above a few bands of 16 diagonals and tiles of 32 cells, every group size."""
    monkeypatch.setenv("DAFS_MCCASKILL_CLUSTER", "1")
    monkeypatch.setenv("DAFS_MCCASKILL_LANES", str(lanes))
    seqs = random_rna(count, n, config_id=21)
    with dafs.McCaskillPlan(seqs, 0.01) as plan:
        assert plan.info()["grid"] == count                      # one CTA per sequence
    assert check_batch(dafs, seqs) <= TOL


def test_banded_kernel_ragged_rescale_and_constraints(dafs, monkeypatch):
    monkeypatch.setenv("DAFS_MCCASKILL_CLUSTER", "1")
    r = random_rna(8, 200, config_id=9)
    seqs = [r[k][:20 + 25 * k] for k in range(8)] + [(b"GGGGCCCCGGGGAAAACCCCGGGGCCCC" * 20)[:500]]
    for th in (0.01, 0.001):
        assert check_batch(dafs, seqs, th) <= TOL
    cseqs = fasta_read(os.path.join(ROOT, "tests", "golden", "RF00017_4.fa"))[1][:3] + random_rna(2, 400, config_id=17)
    cseqs = [q if isinstance(q, bytes) else q.encode() for q in cseqs]
    cons = [constrained_case(q, seed=k) for k, q in enumerate(cseqs)]
    assert check_constrained(dafs, cseqs, cons) <= TOL


def test_banded_kernel_throughput_form_and_repeatable(dafs):
    """160 sequences of 40..120 nt: more sequences than SMs, so the plan takes two CTAs of 384 threads per SM and the
    banded kernel by itself; every sequence against the restatement, and twice the same bits."""
    pool = random_rna(160, 120, config_id=23)
    seqs = [s[:40 + (k % 81)] for k, s in enumerate(pool)]
    with dafs.McCaskillPlan(seqs, 0.01) as plan:
        info = plan.info()
        assert info["block"] == 384 and info["grid"] >= 148
        runs = []
        for rep in range(2):
            plan.run()
            b, lnz = plan.fetch()
            with b:
                runs.append((b.row_ptr.copy(), b.col.copy(), b.val.view(np.uint32).copy(), np.array(lnz)))
        assert all(np.array_equal(x, y) for x, y in zip(*runs))
    assert check_batch(dafs, seqs) <= TOL


VIENNA_GOLD = os.path.join(ROOT, "tests", "golden", "mccaskill_golden.json")


@pytest.mark.skipif(not os.path.exists(VIENNA_GOLD), reason="no ViennaRNA-generated golden file (tools/pin_mccaskill.sh makes it where RNAlib2 2.4.18 exists)")
def test_vienna_golden(dafs):
    """The pin of path (2): RNAfold(bl = true)::calculate of the reference itself, where somebody could run it."""
    import json
    cases = json.load(open(VIENNA_GOLD))["cases"]
    seqs = [c["seq"].encode() for c in cases]
    cons = [None if c["con"] is None else c["con"].encode() for c in cases]
    with dafs.mccaskill_constrained(seqs, cons, 0.01) as b:
        for u, c in enumerate(cases):
            n = len(c["seq"])
            rp, col, val = b.unit(u)
            G = np.zeros((n, n)); G[np.repeat(np.arange(n), np.diff(rp)), col] = val
            R = np.zeros((n, n))
            for i, j, p in c["bp"]:
                R[i, j] = p
            near = (np.abs(G - 0.01) <= TOL) | (np.abs(R - 0.01) <= TOL)
            assert np.abs(G - R)[~near].max(initial=0.0) <= TOL, (u, np.abs(G - R).max())
