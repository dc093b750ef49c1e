"""Breakdown of one end-to-end pair-HMM call (host buffers in, host CSR out).
usage: python tools/prof_e2e.py [torch]   (with "torch": inside a torch CUDA context, like bench.py)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dafs_b200
if len(sys.argv) > 1:
    import torch
    torch.cuda.set_device(0)
    torch.cuda.synchronize()
seqs = dafs_b200.random_rna(256, 300, config_id=4)
pi, pj = dafs_b200.all_pairs(256)
for it in range(5):
    t0 = time.perf_counter()
    p = dafs_b200.PairHmmPlan(seqs, pi, pj, 0.01)
    t1 = time.perf_counter()
    p.run(); ms, nnz = p.sync()
    t2 = time.perf_counter()
    with p.fetch() as b:
        t3 = time.perf_counter()
        n = int(b.nnz)
    t4 = time.perf_counter()
    p.close()
    t5 = time.perf_counter()
    print("create %.1f ms  run+sync %.1f ms (device %.1f)  fetch %.1f ms  free batch %.1f ms  close %.1f ms  total %.1f ms" % (
        (t1 - t0) * 1e3, (t2 - t1) * 1e3, ms, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t5 - t4) * 1e3, (t5 - t0) * 1e3))
