"""Hot spots of an `ncu --page source --csv` export (SASS view).
usage: python tools/ncu_source.py src.csv [top N] [context lines]
Prints the total per stall reason, then the N instructions with most samples."""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
rows = list(csv.reader(open(path)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {h: i for i, h in enumerate(rows[hdr])}
body = [r for r in rows[hdr + 1:] if len(r) > col["# Samples"]]


def num(r, name):
    try:
        return float(r[col[name]])
    except (ValueError, IndexError, KeyError):
        return 0.0


total = sum(num(r, "# Samples") for r in body)
inst = sum(num(r, "Instructions Executed") for r in body)
print("instructions %d  SASS lines %d  samples %d" % (inst, len(body), total))
reasons = [h for h in rows[hdr] if h.startswith("stall_") and "Not Issued" not in h]
tot = {h: sum(num(r, h) for r in body) for h in reasons}
print("stall reasons: " + "  ".join("%s %.1f%%" % (h[6:], 100 * v / max(total, 1)) for h, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v > 0.005 * total))
order = sorted(range(len(body)), key=lambda i: -num(body[i], "# Samples"))[:top]
for i in sorted(order):
    r = body[i]
    why = sorted(((num(r, h), h[6:]) for h in reasons), reverse=True)[:2]
    print("%5d %6.2f%% exec %10d  %-60s %s" % (i, 100 * num(r, "# Samples") / max(total, 1), num(r, "Instructions Executed"),
                                               r[col["Source"]][:60], " ".join("%s=%d" % (n, v) for v, n in why if v)))
