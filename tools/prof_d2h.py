"""Device-to-host copy time of a result-sized page-locked buffer (558 MB), ten times, next to
ten end-to-end pair-HMM calls: is the spread of bench.py's e2e leg the copy?
usage: python tools/prof_d2h.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dafs_b200
torch.cuda.set_device(0)
n = 558 * 1000 * 1000 // 4
d = torch.empty(n, dtype=torch.float32, device="cuda")
h = torch.empty(n, dtype=torch.float32, pin_memory=True)
for rep in range(2):
    ms = []
    for _ in range(10):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        h.copy_(d, non_blocking=True); torch.cuda.synchronize()
        ms.append((time.perf_counter() - t0) * 1e3)
    print("D2H 558 MB (ms):", " ".join("%.1f" % x for x in ms), " -> %.1f GB/s best" % (0.558 / min(ms) * 1e3))
    seqs = dafs_b200.random_rna(256, 300, config_id=4)
    pi, pj = dafs_b200.all_pairs(256)
    ms = []
    for _ in range(11):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
            k = int(b.nnz)
        torch.cuda.synchronize()
        ms.append((time.perf_counter() - t0) * 1e3)
    print("e2e (ms):", " ".join("%.1f" % x for x in ms[1:]))
