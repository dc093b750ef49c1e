"""CPU tests of the N>1 host logic with world_size-2 gloo: the round-robin shard
of bench.py, and the size + max-padded payload all-gather that hands per-rank CSR
fragments to rank 0 -- the same code path bench.py runs over NCCL, on fake
per-rank fragments (no GPU compute here)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_shard_covers_every_pair_once():
    for n, w in [(32640, 8), (2016, 3), (10, 4), (1, 2)]:
        seen = np.concatenate([bench.shard(n, r, w) for r in range(w)])
        assert np.array_equal(np.sort(seen), np.arange(n))
        sizes = [len(bench.shard(n, r, w)) for r in range(w)]
        assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(rank)
    npairs = 11
    mine = bench.shard(npairs, rank, world)
    # a fake per-rank fragment: unit u has (u % 3) + 1 entries valued u + k/10
    frag_nnz = [(int(u) % 3) + 1 for u in mine]
    col = torch.tensor(np.concatenate([np.arange(n) for n in frag_nnz]), dtype=torch.int32)
    val = torch.tensor(np.concatenate([u + np.arange(n) / 10 for u, n in zip(mine, frag_nnz)]), dtype=torch.float32)
    nzo = torch.tensor(np.concatenate([[0], np.cumsum(frag_nnz)]), dtype=torch.int64)
    gathered = bench.gather_csr(dist, world, torch.device("cpu"), col, val, nzo, len(mine))
    # the host-side hand-over: every rank writes its fragment into its region of one shared buffer
    rp = torch.arange(3 * len(mine) + rank, dtype=torch.int32)
    sizes = bench.exchange_sizes(dist, world, torch.device("cpu"), col.numel(), len(mine), rp.numel())
    shared = bench.SharedHostBuffer(dist, rank, world, sizes, pin=False)
    assert shared.fits(sizes)
    shared.put(sizes[rank], nzo, rp, col, val)
    dist.barrier()
    if rank == 0:
        for r in range(world):
            o, p_, c, v = shared.region(r, sizes[r])
            gc, gv, go, gn = gathered[r]
            assert torch.equal(o, go) and torch.equal(c, gc) and torch.equal(v, gv)
            assert torch.equal(p_, torch.arange(3 * gn + r, dtype=torch.int32))
    dist.barrier()
    shared.close()
    if rank == 0:
        out = {}
        for r, (c, v, o, n_units) in enumerate(gathered):
            for k, u in enumerate(bench.shard(npairs, r, world)):
                out[int(u)] = v[o[k]:o[k + 1]].numpy().tolist()
        q.put(out)
    dist.destroy_process_group()


def test_gloo_world2_gather_reassembles_pair_order():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(out) == list(range(11))
    for u, vals in out.items():
        assert np.allclose(vals, [u + k / 10 for k in range((u % 3) + 1)])
