#!/usr/bin/env python
"""Developer tool: per-source-line instruction and stall shares of an ncu report (needs -lineinfo and --import-source on).
    python tools/ncu_lines.py report.ncu-rep [top]"""
import csv, subprocess, sys, io
rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
d = dict(zip(rows[0], rows[2]))
print(d.get("Kernel Name"), "grid", d.get("Grid Size"), "regs", d.get("launch__registers_per_thread"))
for k in ("gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
          "smsp__warps_eligible.avg.per_cycle_active"):
    print(" ", k, d.get(k))
st = {k: float(v) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio") and v not in ("", "n/a")}
print("  stalls per issue:", ", ".join(f"{k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')} {v:.2f}" for k, v in sorted(st.items(), key=lambda x: -x[1])[:7]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
out, cur = [], None
for r in csv.reader(io.StringIO(src)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 8 and r[0].isdigit():
        try:
            out.append((int(r[7]), int(r[4]), cur, r[0], r[1].strip()[:100]))
        except ValueError:
            pass
tot, ts = sum(o[0] for o in out), sum(o[1] for o in out)
for o in sorted(out, reverse=True)[:top]:
    print(f"{o[0] / tot * 100:5.1f}% inst {o[1] / ts * 100:5.1f}% stall  {o[2]}:{o[3]}  {o[4]}")
