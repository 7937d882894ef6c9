#!/usr/bin/env python
"""One line per kernel launch of an .ncu-rep: time, DRAM bytes, registers, achieved occupancy, issue-slot use, top stall."""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
def g(r, k):
    try: return float(r[idx[k]].replace(",", ""))
    except Exception: return float("nan")
stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
UNIT = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
def mb(r, k): return g(r, k) * UNIT.get(rows[1][idx[k]], 1.0)
for r in rows[2:]:
    top = sorted(((g(r, s), s.split("stalled_")[1].split("_per_issue")[0]) for s in stalls), reverse=True)[:3]
    unit = rows[1][idx["gpu__time_duration.sum"]]
    print("%-42s %8.2f %s  dram R/W %7.2f/%6.2f MB  regs %3d  warps %5.1f%%  issue %5.1f%%  fp64 %4.1f%%  stalls %s" % (
        r[idx["Kernel Name"]].split("(")[0][-42:], g(r, "gpu__time_duration.sum"), unit,
        mb(r, "dram__bytes_read.sum"), mb(r, "dram__bytes_write.sum"),
        int(g(r, "launch__registers_per_thread")), g(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), g(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
        " ".join("%s=%.1f" % (n, v) for v, n in top)))
