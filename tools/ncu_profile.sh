#!/bin/bash
# usage: tools/ncu_profile.sh report.ncu-rep out.txt "header line(s)"
# Writes the tracked raw metrics and the SASS hot spots of an ncu report to a text summary.
set -e
rep=$1; out=$2; shift 2
raw=$(mktemp); src=$(mktemp)
ncu -i "$rep" --page raw --csv > "$raw" 2>/dev/null
ncu -i "$rep" --page source --csv > "$src" 2>/dev/null
{
  for h in "$@"; do echo "# $h"; done
  python3 "$(dirname "$0")/ncu_summary.py" "$raw" | grep -E "kernel:|gpu__time_duration|dram__bytes_(read|write).sum  |lts__t_bytes.sum  |launch__registers_per_thread |launch__occupancy_limit|sm__inst_executed.sum.per_cycle_active|sm__inst_executed_pipe_(alu|fma|fp64|lsu|xu).sum.pct_of_peak_sustained_active|smsp__issue_active.avg.pct|sm__warps_active.avg.pct|smsp__thread_inst_executed_per_inst|per_issue_active.ratio" | sed 's/  */ /g'
  echo "# ---- SASS hot spots (ncu --page source) ----"
  python3 "$(dirname "$0")/ncu_source.py" "$src" 25
} > "$out"
python3 - "$raw" <<'PY'
import csv, sys, json
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
def get(name):
    i = hdr.index(name); v = float(vals[i]); u = units[i]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
print(json.dumps({"dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum")}))
PY
rm -f "$raw" "$src"
