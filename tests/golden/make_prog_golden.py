"""Generates tests/golden/prog_golden.json from the REFERENCE's own progressive stage
(oracle/_ref: DAFS::build_tree + DAFS::align lifted from /root/reference/src/dafs.cpp by
oracle/Makefile).  Floats are stored as their IEEE-754 bit patterns.  Run in the build
container:   PYTHONPATH=.:tests python tests/golden/make_prog_golden.py"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_util as ou

cases = []
for name, n, length, seed in [("stemloop4x44", 4, 44, 11), ("stemloop6x56", 6, 56, 12), ("stemloop3x36", 3, 36, 13)]:
    seqs = ou.stemloop_family(n, length, seed)
    lens, mp, bp = ou.pct_inputs(seqs)
    p = ou.ref_pct(lens, mp, bp, 0.25, 0.25)
    mp2, bp2 = ou.pct_to_flat(lens, p)
    r = ou.ref_progressive(lens, mp2, bp2, sim=p["sim"])
    cases.append({
        "name": name, "seqs": [s.decode() for s in seqs],
        "left": r["left"].tolist(), "right": r["right"].tolist(),
        "L": r["L"], "order": r["order"].tolist(),
        "bits": ["".join("1" if b else "0" for b in row) for row in r["bits"]],
        "ss": [int(v) for v in r["ss"].view(np.int32)],
        "log": r["log"].tolist(), "log_score_bits": r["log_score"].view(np.uint32).tolist(),
        "score_bits": int(r["score"].view(np.uint32)),
    })
json.dump({"source": "oracle/_ref ref_progressive (src/dafs.cpp:447-607, 766-951, 1006-1295, 1518-1537; "
                     "-w 4 -t 0.2 -u 0.01 --eta 0.5 -m 600 --no-alifold; inputs relaxed by ref_pct_run -p 0.25 -q 0.25)",
           "cases": cases}, open(os.path.join(HERE, "prog_golden.json"), "w"), indent=0)
print("wrote", len(cases), "cases")
