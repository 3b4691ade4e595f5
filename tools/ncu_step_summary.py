"""Per-kernel totals of an ncu launch list that carries gpu__time_duration.sum, dram__bytes_read.sum and
dram__bytes_write.sum (tools/_run_r2c.sh): count, summed time, share of the step, DRAM bytes per launch and achieved
DRAM GB/s.  Usage: python tools/ncu_step_summary.py in.csv out.txt [hbm_peak_GBps]"""
import csv, json, os, re, sys
from collections import OrderedDict
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = float(sys.argv[3]) if len(sys.argv) > 3 else None
if peak is None:
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
per = OrderedDict()
for r in csv.DictReader(lines):
    name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("void ", "").replace("srb::", "")
    d = per.setdefault(int(r["ID"]), {"name": name})
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    if r["Metric Name"] == "gpu__time_duration.sum":
        d["ns"] = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9, "nsecond": 1, "usecond": 1e3, "msecond": 1e6, "second": 1e9}.get(u, 1)
    elif r["Metric Name"].startswith("dram__bytes"):
        k = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        d["bytes"] = d.get("bytes", 0.0) + v * k
tot = OrderedDict()
for d in per.values():
    a = tot.setdefault(d["name"], [0, 0.0, 0.0])
    a[0] += 1; a[1] += d.get("ns", 0.0); a[2] += d.get("bytes", 0.0)
total = sum(a[1] for a in tot.values())
out = [f"# {len(per)} launches of one eager GAN train step (B = 512), {total / 1e6:.3f} ms summed kernel time "
       f"(ncu, serialised, cold cache; HBM peak used for the last column: {peak:.0f} GB/s)",
       f"{'kernel':58s} {'count':>5s} {'ms':>8s} {'share':>7s} {'MB/launch':>10s} {'GB/s':>8s} {'of HBM':>7s}"]
for n, (c, ns, by) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    gbs = by / ns if ns else 0.0
    out.append(f"{n[:58]:58s} {c:5d} {ns / 1e6:8.3f} {100 * ns / total:6.2f}% {by / c / 1e6:10.2f} {gbs:8.0f} {100 * gbs / peak:6.1f}%")
open(sys.argv[2], "w").write("\n".join(out) + "\n")
print("\n".join(out[:45]))
