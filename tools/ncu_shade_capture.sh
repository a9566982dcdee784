#!/bin/sh
# tools/ncu_shade_capture.sh OUT — one `ncu --set full` capture of k_raygen and the shade kernels of bounces 0 and 1 of ONE batch
# (cornell_dragon 1200x1200, 10 spp, one lane), after the same command has run once without ncu. GPU box only; read the
# report with tools/ncu_summary.py / tools/ncu_lines.py.
set -e
OUT=${1:-gpurun_out/shade_capture}
CMD="python tools/tune_trace.py --variants 4 --lanes 1 --sqrt-spt 1 --reps 0"
$CMD > "$OUT.plain.log" 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_shade|k_raygen' -c 7 -o "$OUT" -f $CMD > "$OUT.ncu.log" 2>&1
ls -la "$OUT.ncu-rep"
