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


def synthetic_bpp(length, family_seed=0, member=0, noise=0.5):
    """A base-pairing-probability-like matrix (dense float32 L x L, only i<j set,
    every kept value > 0.01 like RNAfold::calculate leaves them, fold.cpp:108-114)
    for tests and benchmarks of the DD stage: stems shared by a family plus
    member-specific jitter and low-probability noise pairs.  Stand-in for real
    McCaskill output wherever a test wants many problems cheaply."""
    L = int(length)
    p = np.zeros((L, L), np.float32)
    if L < 8:
        return p
    fam = splitmix64(0xB99000 + 7919 * family_seed, 4 * (L // 8 + 2))
    mem = splitmix64(0xB99000 + 7919 * family_seed + 104729 * (member + 1), 8 * L)
    u = lambda r: float(r >> np.uint64(11)) / float(1 << 53)   # noqa: E731
    rowsum = np.zeros(L)
    m = 0
    for s in range(L // 8 + 1):
        a = int(fam[4 * s] % np.uint64(L))
        span = 5 + int(fam[4 * s + 1] % np.uint64(max(1, L - 5)))
        b = a + span
        n = 3 + int(fam[4 * s + 2] % np.uint64(6))
        base = 0.15 + 0.8 * u(fam[4 * s + 3])
        if u(mem[m]) < 0.15:      # this member lost the stem
            m += 1
            continue
        m += 1
        for t in range(n):
            i, j = a + t, b - t
            if j >= L or j - i < 4:
                break
            v = base * (0.6 + 0.4 * u(mem[m % len(mem)])); m += 1
            v = min(v, 1.0 - rowsum[i], 1.0 - rowsum[j])
            if v > 0.011 and p[i, j] == 0:
                p[i, j] = np.float32(v); rowsum[i] += v; rowsum[j] += v
    nn = int(noise * L)
    for t in range(nn):
        i = int(mem[m % len(mem)] % np.uint64(L)); m += 1
        j = i + 4 + int(mem[m % len(mem)] % np.uint64(L)); m += 1
        if j >= L or p[i, j] != 0:
            continue
        v = 0.011 + 0.3 * u(mem[m % len(mem)]) ** 2; m += 1
        v = min(v, 1.0 - rowsum[i], 1.0 - rowsum[j])
        if v > 0.011:
            p[i, j] = np.float32(v); rowsum[i] += v; rowsum[j] += v
    return p
