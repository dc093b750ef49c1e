"""Small driver for ncu: one DD plan run on a slice of C3 (no timing)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dafs_b200

nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 32
L = int(sys.argv[2]) if len(sys.argv) > 2 else 150
tmax = int(sys.argv[3]) if len(sys.argv) > 3 else 100
seqs = dafs_b200.random_rna(nseq, L, config_id=3)
bpp = [dafs_b200.synthetic_bpp(L, 3 + s % 4, s) for s in range(nseq)]
pi, pj = dafs_b200.all_pairs(nseq)
with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
    pz = [b.dense(u, L) for u in range(len(pi))]
probs = [dict(p_x=bpp[i], p_y=bpp[j], p_z=pz[u], t_max=tmax) for u, (i, j) in enumerate(zip(pi, pj))]
with dafs_b200.DDPlan(probs) as plan:
    plan.run()
    print("ms", plan.sync(), plan.info())
