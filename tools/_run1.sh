python tools/time_e2e.py > gpurun_out/r02_time_e2e.log 2>&1; cat gpurun_out/r02_time_e2e.log
