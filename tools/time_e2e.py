#!/usr/bin/env python
"""tools/time_e2e.py — where the wall-clock time of rtgpu_render(host f64 Buffer) goes on the headline job (GPU box only):
per call, wall clock around the C-ABI call, the library's own ms_total, and the device time of the same job through
rtgpu_render_device (CUDA events), interleaved."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def main():
    import torch
    from rust_raytracer_b200 import api
    from tools.make_dragon_proxy import ensure
    ensure()
    scene, cam, prm = api.load([api.scene_path("cornell_dragon"), "-w=1200", "-t=10", "-s=1000", "--max-depth=20", "--light-bias=0.25"])
    gpu = api.GpuScene(scene, devices=[0])
    host = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
    accum = torch.zeros((cam.height, cam.width, 4), dtype=torch.float32, device="cuda")
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    for i in range(8):
        t0 = time.perf_counter(); gpu.render(cam, prm, out=host); t1 = time.perf_counter()
        st = gpu.stats()
        accum.zero_(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t2 = time.perf_counter(); e0.record(stream); gpu.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream); t3 = time.perf_counter(); e1.record(stream)
        torch.cuda.synchronize(); t4 = time.perf_counter()
        print(json.dumps({"call": i, "render_host_wall_ms": round((t1 - t0) * 1e3, 2), "lib_ms_total": round(st.ms_total, 2),
                          "render_device_events_ms": round(e0.elapsed_time(e1), 2), "render_device_enqueue_wall_ms": round((t3 - t2) * 1e3, 2),
                          "render_device_wall_ms": round((t4 - t2) * 1e3, 2)}), flush=True)


if __name__ == "__main__":
    main()
