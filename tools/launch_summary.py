#!/usr/bin/env python
"""tools/launch_summary.py — per-kernel totals of an ncu launch list.

Input: the CSV of `ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X ...`
Output: one line per kernel name (template arguments kept, parameter list dropped) with launch
count, total / mean / max time and the share of the listed time. Per-launch times under ncu are
cold-cache and serialised: the SHARES are what is compared with bench.py's live numbers.
"""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("void ", "").strip()
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)  # -> microseconds
        rows.append((name, v))
    agg = defaultdict(list)
    for name, v in rows:
        agg[name].append(v)
    total = sum(v for _, v in rows)
    print(f"# {path}: {len(rows)} launches, {total / 1e3:.2f} ms listed")
    print(f"{'kernel':60s} {'n':>5s} {'total ms':>9s} {'mean us':>9s} {'max us':>9s} {'share':>6s}")
    for name, vs in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{name[:60]:60s} {len(vs):5d} {sum(vs) / 1e3:9.2f} {sum(vs) / len(vs):9.1f} {max(vs):9.1f} {sum(vs) / total:6.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
