set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()"
python bench.py > gpurun_out/r02_bench_v10_final.json 2> gpurun_out/bench_v10.err; tail -2 gpurun_out/bench_v10.err; cut -c1-200 gpurun_out/r02_bench_v10_final.json
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/bench_plain_1step.json 2> gpurun_out/bench_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 1250 -c 1165 --csv --log-file gpurun_out/r02_launches_bench_job_final.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_launches.log 2>&1
python tools/launch_summary.py gpurun_out/r02_launches_bench_job_final.csv | head -12
python -c "
import json
d=json.loads(open('gpurun_out/bench_plain_1step.json').read().strip().splitlines()[-1]); print(d['kernel_ms_one_step_profiled'], d['gpu_launches'])"
