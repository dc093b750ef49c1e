"""Generates tests/golden/pairhmm_golden.json from the REFERENCE's own code
(oracle/_ref/libdafs_ref.so = /root/reference sources compiled unmodified).
Run only in the build container:  python tests/golden/make_golden.py
Each case: inputs, nnz, similarity score (dafs.cpp:713-764), and CRC32 of the
kept (row_ptr, col, val) bytes as ProbCons::calculate (align.cpp:60-79) fills
them.  README.md:59 of the reference pins four of the RF00005 scores."""
import json
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_util import ref_pair_dense  # noqa: E402
from dafs_b200.synth import fasta_read, random_rna, homolog_family  # noqa: E402


def csr_of(d, th):
    rows, cols = np.nonzero(d)
    row_ptr = np.zeros(d.shape[0] + 1, np.int32)
    np.add.at(row_ptr, rows + 1, 1)
    row_ptr = np.cumsum(row_ptr).astype(np.int32)
    return row_ptr, cols.astype(np.uint32), d[rows, cols].astype(np.float32)


def case(name, s1, s2, th):
    d, sim = ref_pair_dense(s1, s2, th)
    rp, col, val = csr_of(d, th)
    return {"name": name, "s1": s1.decode(), "s2": s2.decode(), "th": th, "nnz": int(len(col)),
            "sim": float(np.float32(sim)), "crc_row_ptr": zlib.crc32(rp.tobytes()),
            "crc_col": zlib.crc32(col.tobytes()), "crc_val": zlib.crc32(val.tobytes())}


cases = []
names, seqs = fasta_read(os.path.join(HERE, "RF00005_0.fa"))
for i in range(len(seqs)):
    for j in range(i + 1, len(seqs)):
        cases.append(case("RF00005:%d,%d" % (i, j), seqs[i], seqs[j], 0.01))
names, seqs = fasta_read(os.path.join(HERE, "RF00017_4.fa"))
for i, j in [(0, 1), (0, 9), (3, 7), (8, 9)]:
    cases.append(case("RF00017:%d,%d" % (i, j), seqs[i], seqs[j], 0.01))
r = random_rna(6, 150, config_id=3)
cases.append(case("rand150:0,1", r[0], r[1], 0.01))
cases.append(case("rand150:2,3@0.05", r[2], r[3], 0.05))
cases.append(case("rand150:4,5@0.001", r[4], r[5], 0.001))
h = homolog_family(4, 300, config_id=4)
cases.append(case("homolog300:0,1", h[0], h[1], 0.01))
cases.append(case("ragged:1x40", b"G", r[0][:40], 0.01))
cases.append(case("ragged:33x1", r[1][:33], b"a", 0.01))
cases.append(case("ragged:1x1", b"C", b"C", 0.01))
cases.append(case("alphabet", b"acgutnACGUTNxX-acguACGU", b"ACGTTNnnacguRYKMacgtACGU", 0.01))
cases.append(case("wide:20x700", r[0][:20], (r[1] + r[2] + r[3] + r[4] + r[5])[:700], 0.01))
cases.append(case("tall:700x20", (r[1] + r[2] + r[3] + r[4] + r[5])[:700], r[0][:20], 0.01))
json.dump({"source": "oracle/_ref (reference ProbCons::calculate)", "cases": cases},
          open(os.path.join(HERE, "pairhmm_golden.json"), "w"), indent=0)
print(len(cases), "cases")
