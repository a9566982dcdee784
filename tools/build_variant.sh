#!/bin/sh
# tools/build_variant.sh NAME [nvcc defines...] — builds rust_raytracer_b200/_exp/librtgpu_NAME.so with extra -D flags
# (kernel A/B experiments on the GPU box: tools/tune_trace.py --lib rust_raytracer_b200/_exp/librtgpu_NAME.so)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; shift
make -s -C "$ROOT/rust_raytracer_b200/csrc" >/dev/null
mkdir -p "$ROOT/rust_raytracer_b200/_exp" "$ROOT/build/exp"
SRC=$ROOT/rust_raytracer_b200/csrc
/usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v \
  -I"$ROOT/include" -I"$SRC/device" -I"$SRC/host" "$@" -c "$SRC/device/api.cu" -o "$ROOT/build/exp/api_$NAME.o" 2> "$ROOT/build/exp/ptxas_$NAME.log"
OBJS=$(ls "$ROOT"/build/obj/*.o | grep -v '/api.o$')
/usr/local/cuda/bin/nvcc -shared -o "$ROOT/rust_raytracer_b200/_exp/librtgpu_$NAME.so" $OBJS "$ROOT/build/exp/api_$NAME.o" -lz -cudart static
echo "built rust_raytracer_b200/_exp/librtgpu_$NAME.so"
