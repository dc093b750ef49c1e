"""Consistency transformation at benchmark sizes: posteriors from the device, then
dafs_cuda_pct; prints the per-kernel device times.
usage: python tools/prof_pct.py [N] [L] [reps] [--check]   (--check: compare with the oracle; --random: unrelated sequences instead of a family)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import dafs_b200 as db  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
N = int(args[0]) if len(args) > 0 else 256
L = int(args[1]) if len(args) > 1 else 300
reps = int(args[2]) if len(args) > 2 else 3
seqs = db.random_rna(N, L, config_id=4) if "--random" in sys.argv else db.homolog_family(N, L, config_id=4)
lens = np.array([len(s) for s in seqs], np.int32)
mp = db.pairhmm_all(seqs, 0.01)
bp = db.mccaskill_all(seqs, 0.01)
print("N %d L %d  mp nnz %d  bp nnz %d" % (N, L, mp.nnz, bp.nnz), flush=True)
for r in range(reps):
    t0 = time.perf_counter()
    sim, omp, obp = db.pct(lens, mp, bp, 0.25, 0.25)
    t1 = time.perf_counter()
    print("rep %d  host %.1f ms  %s  out mp nnz %d bp nnz %d" % (r, 1e3 * (t1 - t0), db.pct_last_ms(), omp.nnz, obp.nnz), flush=True)
    if r + 1 < reps:
        omp.close(); obp.close()
if "--check" in sys.argv:
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
    import oracle_util as ou
    t0 = time.perf_counter()
    o = ou.oracle_pct(lens, mp, bp, 0.25, 0.25)
    print("oracle %.2f s (1 thread)" % (time.perf_counter() - t0))
    ok = (np.array_equal(sim.view(np.uint32), o["sim"].view(np.uint32))
          and np.array_equal(omp.row_ptr, o["mp_row_ptr"]) and np.array_equal(omp.col, o["mp_col"])
          and np.array_equal(omp.val.view(np.uint32), o["mp_val"].view(np.uint32))
          and np.array_equal(obp.row_ptr, o["bp_row_ptr"]) and np.array_equal(obp.col, o["bp_col"])
          and np.array_equal(obp.val.view(np.uint32), o["bp_val"].view(np.uint32)))
    print("bit-exact with the oracle:", ok)
    sys.exit(0 if ok else 1)
