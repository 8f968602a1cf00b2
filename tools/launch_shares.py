"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.

usage: python tools/launch_shares.py profiles/r02_launches_cfg3.csv > profiles/r02_launch_shares_cfg3.txt"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h, data = rows[hdr], rows[hdr + 1:]
kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in data:
    name = re.sub(r"\(.*", "", r[kn]).replace("void ", "")
    try:
        v = float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[mu], 1e-3)
    except ValueError:
        continue
    agg[name][0] += 1
    agg[name][1] += v
ours = sum(v for n, (c, v) in agg.items() if n.startswith("k_"))
print("# kernel, launches, total us, share of this library's kernel time (cold-cache, serialised launches: compare shares)")
for n, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    share = f"{100 * v / ours:5.1f} %" if n.startswith("k_") else "  (not ours: bench.py's L2 flush)"
    print(f"{n:62s} {c:4d} {v:10.1f} {share}")
