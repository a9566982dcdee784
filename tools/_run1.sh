set -x
timeout 600 python -m pytest tests -m gpu -x -q -k "dropped or broadcast or bounce_parity or converged_image_parity or streamed" > gpurun_out/pytest_gpu_subset.log 2>&1; tail -3 gpurun_out/pytest_gpu_subset.log
python tools/tune_trace.py --variants 4 --ctas 3,4,5,6,8 --sqrt-spt 2 --reps 3 > gpurun_out/r02_tune14_ctas_after_leaf_step.log 2>&1
for sc in 5 6 7; do RTGPU_SHADE_CTAS_PER_SM=$sc python tools/tune_trace.py --variants 4 --ctas 4,5 --sqrt-spt 2 --reps 3 | sed "s/^/shade_ctas=$sc /" >> gpurun_out/r02_tune14_ctas_after_leaf_step.log 2>&1; done
cut -c1-200 gpurun_out/r02_tune14_ctas_after_leaf_step.log
sh tools/ncu_trace_capture.sh gpurun_out/r02_trace_capture_leaf_step; tail -2 gpurun_out/r02_trace_capture_leaf_step.ncu.log
