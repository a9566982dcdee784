#!/usr/bin/env python
"""tools/ncu_summary.py <report.ncu-rep> — key raw metrics + SASS opcode histogram (weighted by executed count)."""
import collections, csv, subprocess, sys, io

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h = r[0]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
for k in keys:
    if k in h:
        i = h.index(k)
        print(f"{k:62s}", [x[i][:40] for x in r[1:]])
for i, k in enumerate(h):
    if 'pcsamp_warps_issue_stalled' in k and 'not_issued' not in k:
        print(f"{k:62s}", [x[i] for x in r[2:]])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = [i for i, x in enumerate(rows) if x and x[0] == 'Address']
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hh = rows[hdr[which]]
body = rows[hdr[which] + 1: hdr[which + 1] - 1 if len(hdr) > which + 1 else None]
iS, iE, iT, iN = hh.index('Source'), hh.index('Instructions Executed'), hh.index('Thread Instructions Executed'), hh.index('# Samples')
op, opT, smp = collections.Counter(), collections.Counter(), collections.Counter()
tot = totT = 0
for x in body:
    if len(x) < len(hh):
        continue
    o = [t for t in x[iS].split() if not t.startswith('@')][0].split('.')[0]
    e, t = int(x[iE] or 0), int(x[iT] or 0)
    op[o] += e; opT[o] += t; smp[o] += int(x[iN] or 0); tot += e; totT += t
print(f"launch #{which}: warp inst {tot}, thread inst {totT}, avg lanes {totT / max(tot, 1):.2f}")
for o, e in op.most_common(24):
    print(f"  {o:10s} {e / tot * 100:5.1f}%  lanes {opT[o] / max(e, 1):5.1f}  samples {smp[o]}")
