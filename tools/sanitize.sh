#!/bin/bash
# compute-sanitizer over the queue / cursor / tail-merging code. ONE tool per gpurun call
# (B200_PROFILING.md: several sanitizer tools in one call have left a GPU unusable): usage
#   tools/sanitize.sh <out.txt> <memcheck|initcheck|racecheck> <cornell|dragon|coverage>
# Small renders with several batches per lane so that chunked RecordCursor reservations, dead-record fill, ping-pong
# class queues and un-reset counter rows across a tail-merging group all run.
OUT=${1:-gpurun_out/r02_sanitizer.txt}
TOOL=${2:-memcheck}
CASE=${3:-cornell}
PY=${PYTHON:-python}
cd "$(dirname "$0")/.."
export RTGPU_LANES=2 RTGPU_BATCH_MPATHS=1 RTGPU_TAIL_GROUP=3
case "$CASE" in
  cornell)  ARGS="assets/scenes/cornell -w=64 -s=256 -t=4" ;;
  dragon)   ARGS="assets/scenes/cornell_dragon -w=128 -s=256 -t=4 --max-depth=20" ;;
  coverage) ARGS="assets/scenes/coverage_test -w=96 -s=64 -t=4" ;;
  *) echo "unknown case $CASE"; exit 2 ;;
esac
echo "==== RTGPU_LANES=2 RTGPU_BATCH_MPATHS=1 RTGPU_TAIL_GROUP=3 compute-sanitizer --tool $TOOL python -m rust_raytracer_b200 $ARGS --gpu" >> "$OUT"
compute-sanitizer --tool "$TOOL" --error-exitcode 9 $PY -m rust_raytracer_b200 $ARGS --gpu --out=/tmp/san.png >> "$OUT" 2>&1
echo "==== exit $?" >> "$OUT"
grep -E "^====|ERROR SUMMARY|Done! Took" "$OUT" | sed 's/^/  /'
