"""CPU tests of the N>1 host logic with world_size-2 gloo: the round-robin shard
of bench.py and the shared host buffer every rank delivers its CSR fragment into
and rank 0 reads in place -- the code path of bench.py's e2e at N > 1, on fake per-rank
fragments (no GPU compute here; the NCCL gather of the library is covered on two GPUs by
tests/test_multi_rank_gpu.py)."""
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
    npairs = 11
    mine = bench.shard(npairs, rank, world)
    # a fake per-rank fragment: unit u has (u % 3) + 1 entries valued u + k/10
    frag_nnz = [(int(u) % 3) + 1 for u in mine]
    col = np.concatenate([np.arange(n) for n in frag_nnz]).astype(np.uint32)
    val = np.concatenate([u + np.arange(n) / 10 for u, n in zip(mine, frag_nnz)]).astype(np.float32)
    nzo = np.concatenate([[0], np.cumsum(frag_nnz)]).astype(np.int64)
    rp = np.arange(3 * len(mine) + rank, dtype=np.int32)
    # the host-side hand-over: every rank writes its fragment into its region of one shared buffer
    # (on a GPU box dafs_cuda_pairhmm_pairs_into does the writing); capacities = maxima over ranks
    t = torch.tensor([len(col), len(mine), len(rp)], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    shared = bench.SharedHostBuffer(dist, rank, world, t.numpy(), pin=False)
    d_nz, d_rp, d_col, d_val = shared.region(rank)
    assert len(d_nz) >= len(nzo) and len(d_col) >= len(col) and len(d_rp) >= len(rp)
    d_nz[:len(nzo)] = nzo; d_rp[:len(rp)] = rp; d_col[:len(col)] = col; d_val[:len(val)] = val
    dist.barrier()
    if rank == 0:
        out = {}
        for r in range(world):
            units = bench.shard(npairs, r, world)
            o, p_, c, v = shared.region(r)
            assert np.array_equal(p_[:3 * len(units) + r], np.arange(3 * len(units) + r))
            for k, u in enumerate(units):
                out[int(u)] = v[o[k]:o[k + 1]].tolist()
                assert np.array_equal(c[o[k]:o[k + 1]], np.arange((int(u) % 3) + 1))
        q.put(out)
    dist.barrier()
    shared.close()
    dist.destroy_process_group()


def test_gloo_world2_shared_host_buffer_reassembles_pair_order():
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
