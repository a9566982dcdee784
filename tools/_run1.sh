L=gpurun_out/r02_tune28_trace_without_sphere_test.log; : > $L
python tools/tune_trace.py --variants 4,41,4,41 --sqrt-spt 5 --reps 3 2>&1 | grep "^{" >> $L
python tools/tune_trace.py --variants 4,41,4,41 --sqrt-spt 2 --reps 3 --profile 1 2>&1 | grep "^{" | sed "s/^/profile /" >> $L
cut -c1-250 $L
