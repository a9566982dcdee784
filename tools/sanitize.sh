#!/bin/bash
# compute-sanitizer over the queue / cursor / tail-merging code: memcheck (out-of-bounds, misaligned),
# initcheck (reads of uninitialised device memory) and racecheck (shared-memory hazards; the kernels use none, kept as a witness).
# Small renders with several batches per lane so that chunked RecordCursor reservations, dead-record fill, ping-pong
# class queues and un-reset counter rows across a tail-merging group all run.  Usage: tools/sanitize.sh <out.txt>
OUT=${1:-gpurun_out/r02_sanitizer.txt}
PY=${PYTHON:-python}
cd "$(dirname "$0")/.."
: > "$OUT"
run() {
  echo "==== $*" >> "$OUT"
  "$@" >> "$OUT" 2>&1
  echo "==== exit $?" >> "$OUT"
}
export RTGPU_LANES=2 RTGPU_BATCH_MPATHS=1 RTGPU_TAIL_GROUP=3
for tool in memcheck initcheck racecheck; do
  run compute-sanitizer --tool $tool --error-exitcode 9 $PY -m rust_raytracer_b200 assets/scenes/cornell -w=64 -s=256 -t=4 --gpu --out=/tmp/san_cornell.png
done
run compute-sanitizer --tool memcheck --error-exitcode 9 $PY -m rust_raytracer_b200 assets/scenes/cornell_dragon -w=128 -s=256 -t=4 --max-depth=20 --gpu --out=/tmp/san_dragon.png
run compute-sanitizer --tool memcheck --error-exitcode 9 $PY -m rust_raytracer_b200 assets/scenes/coverage_test -w=96 -s=64 -t=4 --gpu --out=/tmp/san_cov.png
grep -E "^====|ERROR SUMMARY|Done! Took" "$OUT" | sed 's/^/  /'
