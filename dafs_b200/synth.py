"""Synthetic inputs of the benchmark configs: i.i.d. uniform
ACGU of exact length from splitmix64, seed 0xDAF50000 + 1000*config_id, sequence
s on stream offset s -- so every run, CPU or GPU, sees the same bytes."""
import numpy as np

_M64 = (1 << 64) - 1


def splitmix64(seed, n):
    """n outputs of splitmix64 started at `seed` (vectorised)."""
    with np.errstate(over="ignore"):
        k = (np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
             + np.uint64(seed & _M64))
        z = k
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


_ALPHA = np.frombuffer(b"ACGU", dtype=np.uint8)


def random_rna(nseq, length, config_id=0):
    """nseq random sequences (bytes) of exactly `length`."""
    base = 0xDAF50000 + 1000 * config_id
    out = []
    for s in range(nseq):
        r = splitmix64(base + (s << 32), length)
        out.append(_ALPHA[(r >> np.uint64(62)).astype(np.int64)].tobytes())
    return out


def homolog_family(nseq, length, config_id=0, mutation=0.2):
    """Each sequence = a `mutation`-fraction substituted copy of one root."""
    base = 0xDAF50000 + 1000 * config_id + 500
    root = _ALPHA[(splitmix64(base, length) >> np.uint64(62)).astype(np.int64)].copy()
    out = []
    for s in range(nseq):
        r = splitmix64(base + ((s + 1) << 32), 2 * length)
        seq = root.copy()
        mut = (r[:length] >> np.uint64(11)).astype(np.float64) / float(1 << 53) < mutation
        sub = _ALPHA[(r[length:] >> np.uint64(62)).astype(np.int64)]
        seq[mut] = sub[mut]
        out.append(seq.tobytes())
    return out


def fasta_read(path):
    names, seqs = [], []
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith(">"):
                names.append(line[1:].strip())
                seqs.append("")
            elif line and seqs:
                seqs[-1] += line
    return names, [s.encode("ascii") for s in seqs]
