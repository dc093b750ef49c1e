"""Writes tests/golden/mccaskill_selfgolden.json from oracle/mccaskill_oracle.c
itself -- NOT from the reference (ViennaRNA is absent: parity for this path is
unpinned).  The file only pins the restatement against silent drift."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402
from dafs_b200.synth import fasta_read, random_rna  # noqa: E402

seqs = [s for s in fasta_read(os.path.join(HERE, "RF00005_0.fa"))[1][:3]]
seqs += [fasta_read(os.path.join(HERE, "RF00017_4.fa"))[1][0]]
seqs += [random_rna(1, 150, config_id=3)[0], b"GGGGACGGGGACGGGGACGG"]
cases = []
for s in seqs:
    P, lnz = ou.oracle_mccaskill(s)
    idx = np.argsort(P.ravel())[::-1][:12]
    cases.append({"seq": s.decode(), "lnz": lnz, "top": [[int(k // len(s)), int(k % len(s)), float(P.ravel()[k])] for k in idx]})
json.dump({"source": "oracle/mccaskill_oracle.c (self; parity unpinned)", "cases": cases},
          open(os.path.join(HERE, "mccaskill_selfgolden.json"), "w"), indent=0)
print(len(cases))
