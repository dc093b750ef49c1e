"""Generates tests/golden/pct_golden.json from the REFERENCE's own consistency
transformation (oracle/_ref, compiled from /root/reference by oracle/Makefile).
Floats are stored as their IEEE-754 bit patterns.  Run in the build container:
    PYTHONPATH=.:tests python tests/golden/make_pct_golden.py"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_util as ou
from dafs_b200 import homolog_family, random_rna

cases = []
for name, seqs, wa, ws in [
    ("family5x40", homolog_family(5, 40, config_id=21), 0.25, 0.25),
    ("random4x35", random_rna(4, 35, config_id=22), 0.25, 0.25),
    ("family4x30_neg", homolog_family(4, 30, config_id=23), -1.0, -1.0),
]:
    lens, mp, bp = ou.pct_inputs(seqs)
    r = ou.ref_pct(lens, mp, bp, wa, ws)
    c = {"name": name, "seqs": [s.decode() if isinstance(s, bytes) else s for s in seqs], "w_pct_a": wa, "w_pct_s": ws}
    for k, v in r.items():
        v = np.asarray(v)
        c[k] = (v.view(np.uint32) if v.dtype == np.float32 else v).ravel().tolist()
    cases.append(c)
json.dump({"source": "oracle/_ref ref_pct_run (src/dafs.cpp:155-167, 258-375, 713-764, 1796-1827)", "cases": cases},
          open(os.path.join(HERE, "pct_golden.json"), "w"))
print("wrote", len(cases), "cases")
