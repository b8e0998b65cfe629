"""Per-function share of samples / executed instructions from the source-page CSV written by ncu_summary.py (/tmp/_src.csv)."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1] if len(sys.argv) > 1 else '/tmp/_src.csv')))
srcf = sys.argv[2] if len(sys.argv) > 2 else 'loadbalancedchemistrymodel_b200/csrc/lbchem_kernels_v3.cu'
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Line No"); h = rows[hdr]
isamp = h.index("# Samples"); iinst = h.index("Instructions Executed")
src = open(srcf).read().split('\n')
starts = []
for i, l in enumerate(src, 1):
    m = re.match(r'^(?:static )?__device__ .*?(\w+)\(', l) or re.match(r'^__global__.*? (\w+)\(', l) or re.match(r'^struct (Grp)', l)
    if m: starts.append((i, m.group(1)))
def fn(ln):
    name = 'header'
    for s, n in starts:
        if ln >= s: name = n
        else: break
    return name
def num(x):
    try: return int(float(x))
    except ValueError: return 0
agg = {}
for r in rows[hdr + 1:]:
    if len(r) <= iinst or r[0] == "": continue
    try: ln = int(r[0])
    except ValueError: continue
    a = agg.setdefault(fn(ln), [0, 0]); a[0] += num(r[isamp]); a[1] += num(r[iinst])
ts = sum(a[0] for a in agg.values()) or 1; ti = sum(a[1] for a in agg.values()) or 1
for f, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{f:22s} samples {100*a[0]/ts:5.1f}%  inst {100*a[1]/ti:5.1f}%")
