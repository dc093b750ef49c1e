"""Generates tests/golden/dd_golden.json from the REFERENCE's own code
(oracle/_ref/libdafs_ref.so: nussinov.cpp, needleman_wunsch.cpp and the lifted
DAFS::solve_by_dd of /root/reference, compiled unmodified).  Run only in the
build container:  python tests/golden/make_dd_golden.py
Inputs are regenerated from seeds by tests/oracle_util.dd_problem (p_z comes from
the pair-HMM restatement, itself pinned against the reference); each case keeps
the reference's returned score bits, iteration count, violated count and CRC32
of x, y, z."""
import json
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402

SPECS = [  # name, L, family, l2, n1, n2, t_max, mutation, mix
    ("easy30", 30, 0, None, 1, 1, 600, 0.2, False),
    ("easy73", 73, 1, None, 1, 1, 600, 0.2, False),
    ("profile73", 73, 2, None, 3, 2, 600, 0.2, False),
    ("ragged120x97", 120, 3, 97, 1, 4, 600, 0.2, False),
    ("hard73", 73, 1, None, 1, 1, 600, 0.75, True),
    ("hard150", 150, 4, None, 1, 1, 600, 0.6, True),
    ("hard150b", 150, 5, None, 2, 1, 600, 0.75, True),
    ("hard300", 300, 5, None, 1, 1, 200, 0.75, True),
]
cases = []
for name, L, fam, l2, n1, n2, tmax, mut, mix in SPECS:
    p_x, p_y, p_z = ou.dd_problem(L, family=fam, l2=l2, mutation=mut, mix=mix)
    r = ou.ref_dd(p_x, p_y, p_z, n1, n2, t_max=tmax)
    cases.append({"name": name, "L": L, "family": fam, "l2": l2, "n1": n1, "n2": n2, "t_max": tmax,
                  "mutation": mut, "mix": mix,
                  "score_bits": int(r["score"].view(np.uint32)), "iters": r["iters"],
                  "violated": r["violated"], "crc_x": zlib.crc32(r["x"].tobytes()),
                  "crc_y": zlib.crc32(r["y"].tobytes()), "crc_z": zlib.crc32(r["z"].tobytes())})
    print(name, r["iters"], r["violated"], r["score"])
json.dump({"source": "oracle/_ref (reference solve_by_dd + decoders)", "cases": cases},
          open(os.path.join(HERE, "dd_golden.json"), "w"), indent=0)
