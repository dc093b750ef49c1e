"""bench.py's two legs back to back (device-resident runs, then end-to-end calls) with the SM
clock and power read through NVML after every end-to-end call.  usage: python tools/prof_e2e_bench.py [value_steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dafs_b200
import torch
import pynvml
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
torch.cuda.set_device(0); torch.cuda.synchronize()
seqs = dafs_b200.random_rna(256, 300, config_id=4)
pi, pj = dafs_b200.all_pairs(256)
nvalue = int(sys.argv[1]) if len(sys.argv) > 1 else 13
plan = dafs_b200.PairHmmPlan(seqs, pi, pj, 0.01)
ms = []
for _ in range(nvalue):
    plan.run(); m, _n = plan.sync(); ms.append(m)
plan.close()
print("value leg: " + " ".join("%.1f" % m for m in ms))
for it in range(13):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with dafs_b200.pairhmm_pairs(seqs, pi, pj, 0.01) as b:
        n = int(b.nnz)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) * 1e3
    print("e2e %.1f ms   sm %d MHz  power %.0f W  temp %d C" % (dt, pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
          pynvml.nvmlDeviceGetPowerUsage(h) / 1e3, pynvml.nvmlDeviceGetTemperature(h, pynvml.NVML_TEMPERATURE_GPU)))
