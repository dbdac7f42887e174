"""Turns the ncu outputs brought back in gpurun_out/ into the small tracked summaries under profiles/.

  python tools/summarize_ncu.py <tag> <launches.csv> <full.ncu-rep> <kernel-substring>
writes profiles/<tag>_launches.txt, profiles/<tag>_<kernel>_ncu.txt and profiles/dominant_kernel_traffic.json
"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, launches_csv, rep, kname = sys.argv[1:5]
os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)

with open(launches_csv) as f:
    lines = [ln for ln in f if not ln.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in csv.DictReader(lines):
    try:
        v = float(r["Metric Value"].replace(",", ""))
    except Exception:
        continue
    ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r["Metric Unit"], 1)
    k = re.sub(r"\(.*", "", r["Kernel Name"])[:80]
    agg[k][0] += 1
    agg[k][1] += ns
tot = sum(v[1] for v in agg.values())
with open(os.path.join(ROOT, "profiles", f"{tag}_launches.txt"), "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n# source: {launches_csv}\n")
    for k, (n, ns) in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write(f"{k:82s} launches={n:5d} total_ms={ns / 1e6:10.3f} avg_us={ns / n / 1e3:10.1f} share={ns / tot:.4f}\n")

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keep = re.compile(r"dram__bytes|gpu__time_duration|sm__warps_active|launch__(registers|grid|block|shared|occupancy)|gpu__dram_throughput|"
                  r"sm__throughput|smsp__inst_executed\.sum$|sm__cycles_elapsed\.max|smsp__issue_active|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
                  r"lts__t_bytes\.sum$|smsp__average_warp.*stalled.*ratio|sm__inst_executed_pipe_(fma|alu|lsu|fmaheavy|fmalite)")
out = []
traffic = []
issue = {}
for row in rows[2:]:
    name = row[hdr.index("Kernel Name")]
    if kname not in name:
        continue
    out.append(f"--- {name[:100]}  grid={row[hdr.index('Grid Size')]} block={row[hdr.index('Block Size')]}")
    d = {}
    for i, h in enumerate(hdr):
        if keep.search(h):
            out.append(f"    {h} = {row[i]} {units[i]}")
            d[h] = (row[i], units[i])

    def to_bytes(key):
        v, u = d.get(key, ("0", "byte"))
        return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    traffic.append(to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum"))

    def num(key):
        try:
            return float(d[key][0].replace(",", ""))
        except Exception:
            return None
    issue = {"issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
             "pipe_fma_pct": num("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
             "pipe_alu_pct": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
             "warp_instructions": num("smsp__inst_executed.sum"), "duration_ms": num("gpu__time_duration.sum"),
             "registers_per_thread": num("launch__registers_per_thread")}
with open(os.path.join(ROOT, "profiles", f"{tag}_{kname}_ncu.txt"), "w") as f:
    f.write(f"# ncu --set full --clock-control none --import-source on (source: {rep})\n" + "\n".join(out) + "\n")
if traffic:
    with open(os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json"), "w") as f:
        json.dump({"kernel": kname, "dram_bytes_per_launch": max(traffic), "launches_captured": len(traffic),
                   "source": f"profiles/{tag}_{kname}_ncu.txt", "note": "largest captured launch (a full block of frames)", "issue": issue}, f)
print("ok", len(out), traffic)
