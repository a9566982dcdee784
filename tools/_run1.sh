L=gpurun_out/r02_tune25_no_sphere_no_sky_shade_kernels.log; : > $L
for v in 0 1 0 1; do RTGPU_CONST_TEX_KERNELS=$v python tools/tune_trace.py --variants 4 --sqrt-spt 5 --reps 3 2>&1 | grep "^{" | sed "s/^/specialised=$v /" >> $L; done
for v in 0 1; do RTGPU_CONST_TEX_KERNELS=$v python tools/tune_trace.py --variants 4 --sqrt-spt 2 --reps 3 --profile 1 2>&1 | grep "^{" | sed "s/^/specialised=$v profile /" >> $L; done
for v in 0 1 0 1; do RTGPU_CONST_TEX_KERNELS=$v python tools/tune_trace.py --variants 4 --scene light_test --width 2400 --sqrt-spt 5 --reps 3 2>&1 | grep "^{" | sed "s/^/specialised=$v /" >> $L; done
cut -c1-250 $L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/r02_bench_v9_specialised_shade.json 2> gpurun_out/bench_v9.err; tail -2 gpurun_out/bench_v9.err; cut -c1-200 gpurun_out/r02_bench_v9_specialised_shade.json
