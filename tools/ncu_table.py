"""Compact per-launch table from an `ncu --page raw --csv` export.  Usage: python tools/ncu_table.py file.csv [min_us]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
min_us = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
hdr, data = rows[0], rows[2:]
def col(n): return hdr.index(n) if n in hdr else None
cols = [("us", "gpu__time_duration.sum"), ("grid", "launch__grid_size"), ("smemKB", "launch__shared_mem_per_block_dynamic"),
        ("regs", "launch__registers_per_thread"), ("occ%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"), ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("inst", "smsp__inst_executed.sum"), ("l2hit%", "lts__t_sector_hit_rate.pct")]
stall = [h for h in hdr if "issue_stalled" in h and h.endswith("per_issue_active.ratio") or ("issue_stalled" in h and "ratio" in h)]
print("kernel".ljust(28), *[c[0].rjust(9) for c in cols], " top stalls")
for r in data:
    name = r[col("Kernel Name")].replace("srb::", "").replace("void ", "")
    name = name.split("(")[0][:28]
    try: us = float(r[col("gpu__time_duration.sum")])
    except Exception: continue
    if us < min_us: continue
    vals = []
    for _, h in cols:
        i = col(h)
        v = r[i] if i is not None else "-"
        try: v = f"{float(v.replace(',', '')):.1f}" if "." in v else v
        except Exception: pass
        vals.append(v[:9].rjust(9))
    st = []
    for h in stall:
        try: st.append((float(r[hdr.index(h)]), h.split("issue_stalled_")[1].split("_per")[0].replace(".ratio", "")))
        except Exception: pass
    st.sort(reverse=True)
    print(name.ljust(28), *vals, " ", ", ".join(f"{n}:{v:.1f}" for v, n in st[:4]))
