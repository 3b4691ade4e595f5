"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: a trimmed per-launch CSV (id, kernel, grid,
block, ns) and per-kernel totals with each kernel's share.  Usage: python tools/ncu_launch_summary.py in.csv out_prefix"""
import csv, re, sys
from collections import OrderedDict
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    ns = v * {"ns": 1, "us": 1e3, "usecond": 1e3, "nsecond": 1, "ms": 1e6, "msecond": 1e6, "second": 1e9}.get(unit, 1)
    name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("void ", "").replace("srb::", "")
    rows.append((int(r["ID"]), name, r["Grid Size"], r["Block Size"], ns))
pre = sys.argv[2]
with open(pre + "_launches_trimmed.csv", "w") as f:
    f.write("id,kernel,grid,block,ns\n")
    for r in rows:
        f.write(f'{r[0]},"{r[1][:70]}","{r[2]}","{r[3]}",{r[4]:.0f}\n')
tot = OrderedDict()
for _, n, _, _, ns in rows:
    a = tot.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += ns
total = sum(a[1] for a in tot.values())
with open(pre + "_launches_summary.txt", "w") as f:
    f.write(f"# {len(rows)} launches, {total/1e6:.3f} ms summed kernel time (ncu, serialised, cold cache)\n")
    f.write(f"{'kernel':72s} {'count':>6s} {'ms':>9s} {'share':>7s}\n")
    for n, (c, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{n[:72]:72s} {c:6d} {ns/1e6:9.3f} {100*ns/total:6.2f}%\n")
print(open(pre + "_launches_summary.txt").read()[:3000])
