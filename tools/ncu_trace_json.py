#!/usr/bin/env python
"""tools/ncu_trace_json.py <report.ncu-rep> <rays traversed in the captured launch> [capture note] — writes
profiles/trace_traffic.json from the FIRST k_trace_fast launch of an `ncu --set full` report: DRAM bytes per traversed ray
(bench.py: roofline.traffic) and the issue-roofline inputs (warp instructions, duration, active lanes per instruction)."""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, rays = sys.argv[1], int(sys.argv[2])
note = sys.argv[3] if len(sys.argv) > 3 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units = rows[0], rows[1]
row = [r for r in rows[2:] if "k_trace_fast" in r[h.index("Kernel Name")]][0]


def val(name, want_unit=None):
    i = h.index(name)
    v = float(row[i].replace(",", ""))
    u = units[i]
    scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(u, 1.0)
    return v * scale


rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
out = {
    "kernel": row[h.index("Kernel Name")],
    "grid": int(float(row[h.index("launch__grid_size")])),
    "capture": note,
    "rays_in_launch": rays,
    "duration_ms": val("gpu__time_duration.sum"),
    "inst_executed": val("smsp__inst_executed.sum"),
    "active_lanes": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "achieved_occupancy_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
    "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_ray": (rd + wr) / rays,
}
json.dump(out, open(os.path.join(ROOT, "profiles", "trace_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
