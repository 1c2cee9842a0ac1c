#!/usr/bin/env python
"""Merges the library's per-kernel event times / algorithmic bytes (tools/kernels_a4.py) with the ncu --set full
counters of the same run into one table: achieved GB/s against the measured HBM copy peak, real DRAM bytes, and
the counters that say what limits each kernel.  Usage: ncu_kernel_table.py events.json rep.ncu-rep out.json"""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
events = json.load(open(sys.argv[1]))
rep, out = sys.argv[2], sys.argv[3]
peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
hbm = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
rows = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)))
hdr = rows[0]
M = {"dur": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
     "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "alu": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
     "lsu": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1": "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
     "l2": "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "occ": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
     "waves": "launch__waves_per_multiprocessor"}
units = rows[1]
scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
ncu = {}
for r in rows[2:]:
    d = dict(zip(hdr, r))
    name = d["Kernel Name"].split("(")[0].replace("void ", "").replace("oics::", "")
    e = ncu.setdefault(name, {"n": 0})
    e["n"] += 1
    for k, m in M.items():
        if m in d and d[m] not in ("", "n/a"):
            v = float(d[m].replace(",", "")) * scale.get(units[hdr.index(m)], 1.0)
            e[k] = e.get(k, 0.0) + v
table = []
for ev in events:
    base = ev["kernel"].split("<")[0]
    cand = [n for n in ncu if n.split("<")[0] == base]
    # template arguments: match the channel count / variant where the names carry one
    pick = None
    for n in cand:
        if ev["kernel"] == n or ev["kernel"].replace("NEAREST", "3, 0").replace("LINEAR", "3, 1") == n or "<" not in ev["kernel"]:
            pick = n
            break
    if pick is None and cand:
        pick = cand[0]
    row = {"kernel": ev["kernel"], "launches": ev["launches"], "us_per_launch_events": 1000.0 * ev["ms"] / max(1, ev["launches"])}
    if ev["algorithmic_bytes"] > 0 and ev["ms"] > 0:
        row["algorithmic_bytes_per_launch"] = ev["algorithmic_bytes"] / ev["launches"]
        row["achieved_gbs"] = ev["algorithmic_bytes"] / (ev["ms"] / 1e3) / 1e9
        row["frac_of_hbm_peak"] = row["achieved_gbs"] / hbm
    if pick:
        c, n = ncu[pick], ncu[pick]["n"]
        row["ncu"] = {"kernel": pick, "captured": n, "us_per_launch": c.get("dur", 0) / n,
                      "dram_bytes_per_launch": (c.get("rd", 0) + c.get("wr", 0)) / n}
        for k in ("issue", "alu", "lsu", "l1", "l2", "dram", "occ", "regs", "waves"):
            if k in c:
                row["ncu"][k + ("_pct" if k not in ("regs", "waves") else "")] = round(c[k] / n, 1)
        lim = sorted(((c.get(k, 0) / n, k) for k in ("issue", "alu", "lsu", "l1", "l2", "dram")), reverse=True)
        row["limiter"] = f"{lim[0][1]} {lim[0][0]:.0f} %, {lim[1][1]} {lim[1][0]:.0f} % (occupancy {c.get('occ', 0) / n:.0f} %, {c.get('waves', 0) / n:.1f} waves)"
    table.append(row)
json.dump({"hbm_peak_gbs": hbm, "source": "tools/kernels_a4.py: A4 cards, 8 per batch, device-resident; events = oics_profile, counters = ncu --set full (cold-cache, serialised)",
           "kernels": table}, open(out, "w"), indent=1)
for r in table:
    print(f"{r['kernel']:36s} {r['us_per_launch_events']:9.1f} us  {r.get('achieved_gbs', 0):8.0f} GB/s  {100 * r.get('frac_of_hbm_peak', 0):5.1f} %   {r.get('limiter', '')}")
