"""DD stage on config C3 (N=64, L=150: one 2-leaf problem per pair, 2016 problems,
600-iteration cap) -- device time, iterations/s, and the oracle on a sample.
usage: python tools/bench_dd.py [nseq] [len] [t_max] [mode: random|family] [cpu_sample]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import dafs_b200

nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = int(sys.argv[2]) if len(sys.argv) > 2 else 150
tmax = int(sys.argv[3]) if len(sys.argv) > 3 else 600
mode = sys.argv[4] if len(sys.argv) > 4 else "random"
cpu_n = int(sys.argv[5]) if len(sys.argv) > 5 else 0

if mode == "family":
    seqs = dafs_b200.homolog_family(nseq, L, config_id=3)
    bpp = [dafs_b200.synthetic_bpp(L, 3, s) for s in range(nseq)]
else:
    seqs = dafs_b200.random_rna(nseq, L, config_id=3)
    bpp = [dafs_b200.synthetic_bpp(L, 3 + s % 4, s) for s in range(nseq)]
pi, pj = dafs_b200.all_pairs(nseq)
t0 = time.perf_counter()
with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
    pz = [b.dense(u, L) for u in range(len(pi))]
print("pair-HMM + densify: %.3f s" % (time.perf_counter() - t0))
probs = [dict(p_x=bpp[i], p_y=bpp[j], p_z=pz[u], t_max=tmax) for u, (i, j) in enumerate(zip(pi, pj))]
t0 = time.perf_counter()
plan = dafs_b200.DDPlan(probs)
t_create = time.perf_counter() - t0
for rep in range(3):
    plan.run()
    ms = plan.sync()
t0 = time.perf_counter()
res = plan.fetch()
t_fetch = time.perf_counter() - t0
info = plan.info()
iters = np.array([min(r["iters"] + 1, tmax) for r in res])
print("problems %d  L %d  create %.3f s  device %.3f ms  fetch %.3f s" % (len(probs), L, t_create, ms, t_fetch))
print("info", info)
print("iterations total %d  mean %.1f  capped %d  converged %d  -> %.3f M iter/s, %.1f problems/s, %.2f us/iter/CTA-slot" % (
    iters.sum(), iters.mean(), (iters >= tmax).sum(), sum(r["violated"] == 0 for r in res),
    iters.sum() / ms / 1e3, len(probs) / ms * 1e3, ms * 1e3 * info["grid"] / iters.sum()))
cyc = np.sum([r["cycles"] for r in res], axis=0) / iters.sum()
print("cycles/iter: x fill %.0f tb %.0f | y fill %.0f tb %.0f | z fill %.0f tb %.0f | decode phase %.0f | update phase %.0f" % tuple(cyc))
print("ncbp mean %.1f max %d" % (np.mean([r["ncbp"] for r in res]), max(r["ncbp"] for r in res)))
# how much of the batch time is the queue order?  list-schedule the measured per-problem times
# (decode + update cycles) over the CTA slots in the plan's order (largest cells x min(t_max,
# 1 + ncbp) first), in the ideal order (longest measured time first), and the work bound
import heapq
dur = np.array([r["cycles"][6] + r["cycles"][7] for r in res]) / 1.965e6          # ms at 1965 MHz
ncbp = np.array([r["ncbp"] for r in res])
def makespan(order):
    slots = [0.0] * info["grid"]
    heapq.heapify(slots)
    for k in order:
        heapq.heappush(slots, heapq.heappop(slots) + dur[k])
    return max(slots)
plan_order = sorted(range(len(res)), key=lambda k: -min(tmax, 1 + ncbp[k]))
print("queue model (ms): plan order %.1f | longest-first %.1f | work / slots %.1f | longest problem %.1f | measured %.1f" % (
    makespan(plan_order), makespan(np.argsort(-dur)), dur.sum() / info["grid"], dur.max(), ms))
cap = iters >= tmax
rank = {k: r for r, k in enumerate(plan_order)}
print("capped problems: %d, queue positions min %d median %d max %d of %d (slots %d)" % (
    cap.sum(), min(rank[k] for k in np.where(cap)[0]) if cap.any() else -1,
    int(np.median([rank[k] for k in np.where(cap)[0]])) if cap.any() else -1,
    max(rank[k] for k in np.where(cap)[0]) if cap.any() else -1, len(res), info["grid"]))
if cpu_n:
    import oracle_util as ou
    t0 = time.perf_counter(); it = 0
    for u in range(0, len(probs), max(1, len(probs) // cpu_n)):
        o = ou.oracle_dd(probs[u]["p_x"], probs[u]["p_y"], probs[u]["p_z"], t_max=tmax)
        g = res[u]
        assert o["iters"] == g["iters"] and o["score"].view(np.uint32) == g["score"].view(np.uint32), u
        assert all(np.array_equal(o[k], g[k]) for k in "xyz")
        it += min(o["iters"] + 1, tmax)
    dt = time.perf_counter() - t0
    print("oracle (1 core): %d iterations in %.2f s -> %.1f us/iter ; GPU speed-up per core %.0fx" % (
        it, dt, dt / it * 1e6, (iters.sum() / (ms * 1e-3)) / (it / dt)))
plan.close()
if os.environ.get("DAFS_BENCH_DD_DUMP"):
    np.save(os.environ["DAFS_BENCH_DD_DUMP"], np.array([[r["ncbp"], min(r["iters"] + 1, tmax), r["violated"], dur[k] * 1e3] for k, r in enumerate(res)]))
