#!/usr/bin/env python
"""Run under torchrun with N >= 2 ranks (one per GPU): every rank computes its round-robin shard
of a small all-pairs batch; the fragments reach rank 0 both ways bench.py uses --
  (a) dafs_cuda_pairhmm_plan_gather: the library's NCCL gather-v to rank 0's GPU,
  (b) dafs_cuda_pairhmm_pairs_into: per-rank delivery into shared host memory --
and rank 0 compares both, reassembled in pair order, with the one-rank CSR BYTE FOR BYTE
(SURVEY.md section 4 (iv)).  Prints "multirank ok ..." on success; any mismatch raises.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/multirank_check.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import bench
    import dafs_b200
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    seqs = [s[:120 + 5 * k] for k, s in enumerate(dafs_b200.random_rna(24, 300, config_id=6))]
    pi, pj = dafs_b200.all_pairs(len(seqs))
    mine = bench.shard(len(pi), rank, world)
    mpi, mpj = pi[mine], pj[mine]
    # (a) device gather
    uid = [dafs_b200.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    comm = dafs_b200.Comm(uid[0], rank, world)
    plan = dafs_b200.PairHmmPlan(seqs, mpi, mpj, 0.01)
    plan.run()
    _, nnz_mine = plan.sync()
    frags_dev = None
    for _ in range(2):                                    # twice: buffers of the first gather are reused
        g = comm.gather(plan, 0)
        if rank == 0:
            frags_dev = [tuple(a.copy() for a in f) for f in g.to_host()]
        else:
            assert g.empty
        g.close()
    # (b) host delivery into the shared buffer
    rows_mine = int(sum(len(seqs[i]) + 1 for i in mpi))
    t = torch.tensor([nnz_mine, len(mine), rows_mine], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    shared = bench.SharedHostBuffer(dist, rank, world, t.cpu().numpy())
    dst = dafs_b200.HostCsr(*shared.region(rank))
    assert dafs_b200.pairhmm_pairs_into(seqs, mpi, mpj, 0.01, dst) == nnz_mine
    # a too-small destination is refused with the needed size
    small = dafs_b200.HostCsr(np.zeros(len(mine) + 1, np.int64), np.zeros(rows_mine, np.int32),
                              np.zeros(8, np.uint32), np.zeros(8, np.float32))
    try:
        dafs_b200.pairhmm_pairs_into(seqs, mpi, mpj, 0.01, small)
        raise AssertionError("expected DAFS_CUDA_ENOMEM")
    except dafs_b200.DafsCudaError as e:
        assert e.code == -4 and small.nnz == nnz_mine
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as one:
            for name, frags in (("nccl gather", frags_dev),
                                ("shared host", [None] * world)):
                total = 0
                for r in range(world):
                    units = bench.shard(len(pi), r, world)
                    if frags[r] is None:
                        nz, rp, col, val = shared.region(r)
                    else:
                        nz, rp, col, val = frags[r]
                    assert nz[0] == 0
                    row0 = 0
                    for k, u in enumerate(units):
                        orp, ocol, oval = one.unit(int(u))
                        z0, z1 = int(nz[k]), int(nz[k + 1])
                        assert rp[row0:row0 + len(orp)].tobytes() == orp.tobytes(), (name, r, k)
                        assert col[z0:z1].tobytes() == ocol.tobytes(), (name, r, k)
                        assert val[z0:z1].tobytes() == oval.tobytes(), (name, r, k)
                        row0 += len(orp)
                        total += z1 - z0
                assert total == one.nnz, (name, total, one.nnz)
            print("multirank ok: %d ranks, %d pairs, %d posteriors identical to the one-rank CSR "
                  "(NCCL %d gather-v and shared host delivery)" % (world, len(pi), one.nnz, comm.nccl_version()))
    dist.barrier()
    shared.close()
    plan.close()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
