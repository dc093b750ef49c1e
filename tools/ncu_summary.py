"""Prints the metrics we track from an `ncu --page raw --csv` dump."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
pat = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
       'sm__inst_executed.sum', 'smsp__inst_executed.avg.per_cycle_active', 'sm__inst_executed_pipe_',
       'smsp__issue_active.avg.pct', 'lts__t_bytes.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
       'smsp__average_warp', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct',
       'smsp__thread_inst_executed_per_inst_executed', 'smsp__average_warps_issue_stalled', 'launch__occupancy_limit',
       'sm__pipe_', 'smsp__pcsamp_warps_issue_stalled', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared']
for vals in rows[2:]:
    print("== kernel:", vals[hdr.index("Kernel Name")][:60] if "Kernel Name" in hdr else "")
    for h, u, v in zip(hdr, units, vals):
        if any(w in h for w in pat) and v not in ("", "0", "n/a"):
            print("%-90s %-14s %s" % (h, u, v))
