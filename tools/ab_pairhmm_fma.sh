#!/bin/bash
# A/B: what bit-exactness costs the pair-HMM kernel.  Builds a second library whose LOOKUP polynomial is fused
# (3 FMA instead of 3 x (MUL, ADD): 11 instead of 14 instructions per LOG_ADD), runs C4 with both and compares the
# posteriors.  The fused build is never shipped: north_star asks for 1e-4, the parity tests for equal bits.
# usage (on a GPU box): tools/ab_pairhmm_fma.sh
set -e
cd "$(dirname "$0")/.."
mkdir -p exp
(cd dafs_b200/csrc && nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-O2,-pthread,-ffp-contract=off \
   --fmad=false -DDAFS_PAIRHMM_FMA -I../../include -c pairhmm.cu -o /tmp/pairhmm_fma.o)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o exp/libdafs_fma.so /tmp/pairhmm_fma.o dafs_b200/csrc/common.o dafs_b200/csrc/dd.o \
   dafs_b200/csrc/mccaskill.o dafs_b200/csrc/pct.o dafs_b200/csrc/prog.o dafs_b200/csrc/comm.o -lcudart_static -lpthread -ldl -lrt
python - <<'PY'
import os, subprocess, sys, json
import numpy as np
code = r'''
import os, sys, numpy as np
sys.path.insert(0, os.getcwd())
import dafs_b200 as db
seqs = db.random_rna(256, 300, config_id=4)
pi, pj = db.all_pairs(256)
with db.PairHmmPlan(seqs, pi, pj, 0.01) as plan:
    for _ in range(4):
        plan.run(); ms, nnz = plan.sync()
    b = plan.fetch()
    np.savez(sys.argv[1], ms=ms, nnz=nnz, row_ptr=b.row_ptr, col=b.col, val=b.val, nz_off=b.nz_off)
    b.close()
'''
out = {}
for name, lib in (("exact", "dafs_b200/libdafs_cuda.so"), ("fma", "exp/libdafs_fma.so")):
    env = dict(os.environ, DAFS_CUDA_LIB=os.path.join(os.getcwd(), lib))
    subprocess.run([sys.executable, "-c", code, "/tmp/ab_%s.npz" % name], check=True, env=env)
    out[name] = np.load("/tmp/ab_%s.npz" % name)
a, f = out["exact"], out["fma"]
res = {"ms_exact": float(a["ms"]), "ms_fma": float(f["ms"]), "nnz_exact": int(a["nnz"]), "nnz_fma": int(f["nnz"])}
if res["nnz_exact"] == res["nnz_fma"] and np.array_equal(a["col"], f["col"]) and np.array_equal(a["row_ptr"], f["row_ptr"]):
    d = np.abs(a["val"] - f["val"])
    res.update(same_pattern=True, max_abs_diff=float(d.max()), mean_abs_diff=float(d.mean()), differing_values=int((a["val"].view(np.uint32) != f["val"].view(np.uint32)).sum()))
else:
    res.update(same_pattern=False, entries_only_in_one=abs(res["nnz_exact"] - res["nnz_fma"]))
print(json.dumps(res))
PY
