"""How much of the progressive stage is the wave barrier?  Runs the device pipeline and compares
the sum over waves of the slowest node (what wave-synchronous scheduling pays) with the longest
root-to-leaf chain of node costs (the bound of dependency-driven scheduling), in DD iterations.
usage: python tools/prog_critical_path.py nseq len family|random|file.fa"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dafs_b200 as db
nseq, L, mode = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
seqs = db.fasta_read(mode)[1] if mode.endswith(".fa") else (db.homolog_family(nseq, L, config_id=4) if mode == "family" else db.random_rna(nseq, L, config_id=4))
lens = np.array([len(s) for s in seqs], np.int32)
n = len(seqs)
with db.pairhmm_all(seqs, 0.01) as mp, db.mccaskill_all(seqs, 0.01) as bp:
    sim, mp2, bp2 = db.pct(lens, mp, bp, 0.25, 0.25)
    left, right, _ = db.guide_tree(sim)
    g = db.progressive(lens, mp2, bp2, left, right)
it = np.minimum(g["node_iters"] + 1, 600).astype(np.int64)
it[:n] = 0
cost = it * g["node_L"].astype(np.int64)          # iterations x columns ~ time of the latency-bound solve
level = np.zeros(2 * n - 1, int); chain = np.zeros(2 * n - 1, np.int64); chain_it = np.zeros(2 * n - 1, np.int64)
for k in range(n, 2 * n - 1):
    level[k] = 1 + max(level[left[k]], level[right[k]])
    chain[k] = cost[k] + max(chain[left[k]], chain[right[k]])
    chain_it[k] = it[k] + max(chain_it[left[k]], chain_it[right[k]])
waves = sum(cost[(level == lv)].max() for lv in range(1, level.max() + 1))
waves_it = sum(it[(level == lv)].max() for lv in range(1, level.max() + 1))
print("nodes %d, waves %d, DD iterations total %d" % (n - 1, level.max(), it.sum()))
print("wave-synchronous: sum of per-wave maxima = %d iterations (%d iteration-columns)" % (waves_it, waves))
print("dependency-driven bound: longest chain   = %d iterations (%d iteration-columns) -> %.2fx" % (chain_it[-1], chain[-1], waves / max(chain[-1], 1)))
print("device ms: DD %.1f, total %.1f" % (g["ms"]["dd"], g["ms"]["total"]))
