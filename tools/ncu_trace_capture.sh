#!/bin/sh
# tools/ncu_trace_capture.sh OUT — one `ncu --set full` capture of bounce 1 of ONE batch (cornell_dragon 1200x1200, 10 spp, one lane):
# k_trace_fast and the three shade kernels, after the same command has run once without ncu. GPU box only; read the report with
# tools/ncu_summary.py / tools/ncu_lines.py / tools/ncu_trace_json.py (library = the build the report was taken from).
set -e
OUT=${1:-gpurun_out/trace_capture}
CMD="python tools/tune_trace.py --variants 4 --lanes 1 --sqrt-spt 1 --reps 0"
$CMD > "$OUT.plain.log" 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_trace_fast|k_shade' -s 4 -c 4 -o "$OUT" -f $CMD > "$OUT.ncu.log" 2>&1
ls -la "$OUT.ncu-rep"
