"""End-to-end device pipeline on one input set: McCaskill -> pair-HMM -> consistency
transformation -> guide tree -> progressive alignment (all through the C ABI), with the
wall time of every stage; optionally the restated reference loop (oracle) on the first
`cpu_n` sequences of the same set for a CPU time of the progressive stage.
usage: python tools/bench_prog.py [nseq] [len] [mode: family|random] [t_max] [cpu_n] [reps]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dafs_b200 as db

nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = int(sys.argv[2]) if len(sys.argv) > 2 else 150
mode = sys.argv[3] if len(sys.argv) > 3 else "family"
tmax = int(sys.argv[4]) if len(sys.argv) > 4 else 600
cpu_n = int(sys.argv[5]) if len(sys.argv) > 5 else 0
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 2

if mode.endswith(".fa"):                       # a FASTA file instead of synthetic input (nseq / len ignored)
    seqs = db.fasta_read(mode)[1]
    nseq, L = len(seqs), max(len(q) for q in seqs)
else:
    seqs = db.homolog_family(nseq, L, config_id=4) if mode == "family" else db.random_rna(nseq, L, config_id=4)
lens = np.array([len(s) for s in seqs], np.int32)
db.init(1)
out = {"nseq": nseq, "len": L, "mode": mode, "t_max": tmax}
for rep in range(reps):
    t = [time.perf_counter()]
    bp = db.mccaskill_all(seqs, 0.01); t.append(time.perf_counter())
    mp = db.pairhmm_all(seqs, 0.01); t.append(time.perf_counter())
    sim, mp2, bp2 = db.pct(lens, mp, bp, 0.25, 0.25); t.append(time.perf_counter())
    left, right, _ = db.guide_tree(sim); t.append(time.perf_counter())
    g = db.progressive(lens, mp2, bp2, left, right, t_max=tmax); t.append(time.perf_counter())
    st = dict(zip(("mccaskill", "pairhmm", "pct", "guide_tree", "progressive"), [(b - a) * 1e3 for a, b in zip(t, t[1:])]))
    st["total"] = (t[-1] - t[0]) * 1e3
    print("rep %d: " % rep + "  ".join("%s %.1f ms" % kv for kv in st.items()))
    print("   progressive: waves %d, L %d, DD iterations %d, launches %d, device ms: averaging %.2f, DD %.2f; DD plans wall %.1f ms"
          % (g["waves"], g["L"], g["total_iters"], g["gpu_launches"], g["ms"]["average"], g["ms"]["dd"], g["ms"]["dd_host"]))
    out["stages_ms"] = st
    out["progressive"] = {"waves": g["waves"], "L": g["L"], "total_iters": g["total_iters"], "launches": g["gpu_launches"],
                          "ms": g["ms"], "score": float(g["score"]), "posteriors_in": int(mp2.nnz)}
    if rep + 1 < reps:
        for b in (bp, mp, mp2, bp2):
            b.close()
if cpu_n >= 2:
    import oracle_util as ou
    sub = seqs[:cpu_n]
    sl = lens[:cpu_n]
    with db.pairhmm_all(sub, 0.01) as m, db.mccaskill_all(sub, 0.01) as b:
        s2, m2, b2 = db.pct(sl, m, b, 0.25, 0.25)
        l2, r2, _ = db.guide_tree(s2)
        t0 = time.perf_counter()
        gg = db.progressive(sl, m2, b2, l2, r2, t_max=tmax)
        t_gpu = time.perf_counter() - t0
        t0 = time.perf_counter()
        fn = ou.ref_progressive if ou.have_ref() else None
        o = ou.ref_progressive(sl, m2, b2, tree=(l2, r2), t_max=tmax) if fn else ou.oracle_progressive(sl, m2, b2, (l2, r2), t_max=tmax)
        t_cpu = time.perf_counter() - t0
        same = (o["L"] == gg["L"] and np.array_equal(o["bits"], gg["bits"]) and np.array_equal(o["ss"], gg["ss"])
                and o["score"].view(np.uint32) == gg["score"].view(np.uint32))
        print("first %d sequences: progressive on the device %.1f ms, %s on one core %.1f ms (%.1fx); identical: %s"
              % (cpu_n, t_gpu * 1e3, "reference (oracle/_ref)" if fn else "restated reference", t_cpu * 1e3, t_cpu / t_gpu, same))
        out["cpu"] = {"nseq": cpu_n, "kind": "reference" if fn else "port", "cores": 1, "ms": t_cpu * 1e3, "device_ms": t_gpu * 1e3,
                      "identical": bool(same)}
print(json.dumps(out))
