"""End-to-end pair-HMM call the way bench.py's e2e leg makes it: create, run, fetch (no sync in
between, so fetch copies finished chunks under the running kernel).  usage: python tools/prof_e2e_overlap.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dafs_b200
import torch
torch.cuda.set_device(0); torch.cuda.synchronize()
seqs = dafs_b200.random_rna(256, 300, config_id=4)
pi, pj = dafs_b200.all_pairs(256)
one_call = len(sys.argv) > 1 and sys.argv[1] == "pairs"
for it in range(16):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if one_call:
        with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
            n = int(b.nnz)
        torch.cuda.synchronize()
        print("pairhmm_pairs total %.1f ms  nnz %d" % ((time.perf_counter() - t0) * 1e3, n))
        continue
    p = dafs_b200.PairHmmPlan(seqs, pi, pj, 0.01)
    t1 = time.perf_counter()
    p.run()
    with p.fetch() as b:
        t3 = time.perf_counter()
        n = int(b.nnz)
    p.close()
    t5 = time.perf_counter()
    print("chunks %s  create %.1f ms  run+fetch %.1f ms  total %.1f ms  nnz %d" % (os.environ.get("DAFS_PAIRHMM_CHUNKS", "auto"), (t1 - t0) * 1e3, (t3 - t1) * 1e3, (t5 - t0) * 1e3, n))
