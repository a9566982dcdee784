L=gpurun_out/r02_tune23_const_texture_shade_kernels.log; : > $L
for v in 0 1 0 1; do RTGPU_CONST_TEX_KERNELS=$v python tools/tune_trace.py --variants 4 --sqrt-spt 5 --reps 3 2>&1 | grep "^{" | sed "s/^/ctex=$v /" >> $L; done
for v in 0 1; do RTGPU_CONST_TEX_KERNELS=$v python tools/tune_trace.py --variants 4 --sqrt-spt 2 --reps 3 --profile 1 2>&1 | grep "^{" | sed "s/^/ctex=$v profile /" >> $L; done
for sc in "cornell 600 3" "light_test 2400 5"; do set -- $sc; for v in 0 1 0 1; do RTGPU_CONST_TEX_KERNELS=$v python tools/tune_trace.py --variants 4 --scene $1 --width $2 --sqrt-spt $3 --reps 3 2>&1 | grep "^{" | sed "s/^/ctex=$v /" >> $L; done; done
cut -c1-250 $L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
