python tools/tune_trace.py --variants 4,38,39,40,4,38 --sqrt-spt 5 --reps 2 > gpurun_out/r02_tune19_leaf_step_ray_prefetch.log 2>&1
python tools/tune_trace.py --variants 4,38,39,40 --sqrt-spt 2 --reps 3 --profile 1 >> gpurun_out/r02_tune19_leaf_step_ray_prefetch.log 2>&1
cut -c1-200 gpurun_out/r02_tune19_leaf_step_ray_prefetch.log
