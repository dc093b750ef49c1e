"""McCaskill stage: device time for N sequences of length L, oracle on a sample.
usage: python tools/bench_mc.py [nseq] [len] [cpu_sample]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import dafs_b200

nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 256
L = int(sys.argv[2]) if len(sys.argv) > 2 else 300
cpu_n = int(sys.argv[3]) if len(sys.argv) > 3 else 0
seqs = dafs_b200.random_rna(nseq, L, config_id=4)
t0 = time.perf_counter()
plan = dafs_b200.McCaskillPlan(seqs, 0.01)
t_create = time.perf_counter() - t0
for rep in range(3):
    plan.run()
    ms, nnz = plan.sync()
info = plan.info()
flop = nseq * (3.5 * L ** 3 + 1100.0 * L ** 2)
print("sequences %d  L %d  create %.3f s  device %.3f ms  nnz %d  -> %.1f seq/s, ~%.2f TFLOP/s FP64 (SURVEY estimate 3.5 L^3 + 1100 L^2)" % (
    nseq, L, t_create, ms, nnz, nseq / ms * 1e3, flop / ms / 1e9))
print("info", info)
if cpu_n:
    import oracle_util as ou
    t0 = time.perf_counter()
    for s in seqs[:cpu_n]:
        ou.oracle_mccaskill(s)
    dt = (time.perf_counter() - t0) / cpu_n
    print("oracle (1 core, FP64, two inside passes): %.3f s/seq -> GPU speed-up per core %.0fx" % (dt, dt / (ms * 1e-3 / nseq)))
plan.close()
