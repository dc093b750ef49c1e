"""Aggregate device-to-host bandwidth of N ranks (torchrun), each copying its share of a C4-sized result
(558 MB / N) from its own GPU: (a) into its own cudaHostAlloc buffer, (b) into its region of one POSIX
shared-memory buffer that it page-locks (bench.SharedHostBuffer), pages first touched by rank 0,
(c) the same with every rank bound to the NUMA node of its GPU and touching its own region first.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/prof_d2h_ranks.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dist.init_process_group("nccl" if world > 1 else "gloo", device_id=torch.device("cuda", lr) if world > 1 else None)
import bench

node, cpus = bench.gpu_numa_node(lr)
print("rank %d gpu %d numa node %s cpus %s affinity %d" % (rank, lr, node, (str(len(cpus)) if cpus else "-"), len(os.sched_getaffinity(0))), flush=True)
nbytes = 558 * 1000 * 1000 // world // 4096 * 4096
d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")


def timed(dst, label):
    ms = []
    for _ in range(6):
        dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
        dst.copy_(d, non_blocking=True)
        torch.cuda.synchronize(); dist.barrier()
        ms.append((time.perf_counter() - t0) * 1e3)
    if rank == 0:
        print("%-64s ms: %s -> %.1f GB/s aggregate (best)" % (label, " ".join("%.2f" % x for x in ms), nbytes * world / min(ms[1:]) / 1e6), flush=True)


h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
timed(h, "(a) private cudaHostAlloc per rank")
for mode, label in [(False, "(b) shared buffer, first touch by rank 0"), (True, "(c) shared buffer, NUMA-local first touch")]:
    sh = bench.SharedHostBuffer(dist, rank, world, (nbytes // 8, 16, 16), numa_local=mode, gpu=lr, touch_all_by_rank0=not mode)
    view = torch.from_numpy(sh.bytes[rank * sh.stride:rank * sh.stride + nbytes])
    timed(view, label + (" [pinned]" if sh.pinned else " [NOT pinned]"))
    del view
    sh.close()
    dist.barrier()
if bench.bind_to_gpu_node(lr):
    h2 = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    timed(h2, "(d) private cudaHostAlloc per rank, process bound to the GPU's node")
dist.destroy_process_group()
