"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))


def toint(x):
    try:
        return int(float(x))
    except ValueError:
        return 0

top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
h = rows[hdr]
isamp = h.index("# Samples"); iinst = h.index("Instructions Executed")
stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
data = []
for r in rows[hdr + 1:]:
    if len(r) <= iinst or r[0] == "":
        continue       # SASS rows have an empty line number; line rows carry the per-line totals
    try:
        ln = int(r[0])
    except ValueError:
        continue
    st = {c: toint(r[i]) for i, c in stall_cols}
    data.append((ln, r[1], toint(r[isamp]), toint(r[iinst]), st))
tot = sum(d[2] for d in data) or 1; toti = sum(d[3] for d in data) or 1
print("total samples", tot, "warp instructions", toti)
agg = {}
for d in data:
    for k, v in d[4].items():
        agg[k] = agg.get(k, 0) + v
print("stalls:", {k: round(100 * v / tot, 1) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for d in sorted(data, key=lambda d: -d[2])[:top]:
    s = sorted(d[4].items(), key=lambda kv: -kv[1])[:2]
    print(f"{d[0]:5d} {100*d[2]/tot:5.1f}% samp {100*d[3]/toti:5.1f}% inst {s[0][0][6:]:>10s}/{s[1][0][6:]:<10s} {d[1].strip()[:100]}")
