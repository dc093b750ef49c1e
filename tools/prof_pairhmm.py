"""Small driver for ncu: runs the C4 pair-HMM plan a few times (no timing)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dafs_b200

nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 256
length = int(sys.argv[2]) if len(sys.argv) > 2 else 300
runs = int(sys.argv[3]) if len(sys.argv) > 3 else 2
seqs = dafs_b200.random_rna(nseq, length, config_id=4)
pi, pj = dafs_b200.all_pairs(nseq)
with dafs_b200.PairHmmPlan(seqs, pi, pj, 0.01) as plan:
    for _ in range(runs):
        plan.run()
        ms, nnz = plan.sync()
    print("pairs", len(pi), "nnz", nnz, "ms", ms, plan.info())
