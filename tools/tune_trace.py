#!/usr/bin/env python
"""tools/tune_trace.py — A/B the trace-kernel variants on the bench workload (GPU box only).

For every RTGPU_TRACE_VARIANT in --variants (and every CTAs/SM in --ctas) renders the same
10-spp slice of cornell_dragon 1200x1200 and prints per-family kernel time, Mrays/s and the
image difference against variant 0 (the v1 one-ray-per-thread kernel), so that a wrong-but-fast
variant cannot be picked.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", default="0,1,2,3,4,5,6,7,8,9")
    ap.add_argument("--ctas", default="4")
    ap.add_argument("--lanes", default="4")
    ap.add_argument("--batch", default="40", help="RTGPU_BATCH_MPATHS values")
    ap.add_argument("--profile", type=int, default=0, help="1: per-kernel events (forces one lane)")
    ap.add_argument("--scene", default="cornell_dragon")
    ap.add_argument("--width", type=int, default=1200)
    ap.add_argument("--sqrt-spt", type=int, default=1)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--max-depth", type=int, default=20)
    ap.add_argument("--lib", default="", help="alternative librtgpu.so (tools/build_variant.sh)")
    args = ap.parse_args()
    import torch
    from rust_raytracer_b200 import api, _abi as abi
    from tools.make_dragon_proxy import ensure
    if args.lib:
        api.LIB_PATH = os.path.abspath(args.lib)
    ensure()
    T, k = 10, args.sqrt_spt
    scene_args = [] if args.scene == "default" else [api.scene_path(args.scene)]   # "default" = built-in golden_monkey scene
    scene, cam, prm = api.load(scene_args + [f"-w={args.width}", f"-t={T}", f"-s={T * k * k}", f"--max-depth={args.max_depth}", "--light-bias=0.25"])
    if args.profile:
        prm.flags |= abi.FLAG_PROFILE
    gpu = api.GpuScene(scene, devices=[0])
    W, H = cam.width, cam.height
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    stream = torch.cuda.Stream()  # a real stream: handle 0 would make the library fall back to its own
    torch.cuda.set_stream(stream)
    ref = None
    for v in [int(x) for x in args.variants.split(",")]:
        for ctas, lanes, batch in [(int(x), int(y), int(z)) for x in args.ctas.split(",") for y in args.lanes.split(",")
                                   for z in args.batch.split(",")]:
            os.environ["RTGPU_TRACE_VARIANT"] = str(v)
            os.environ["RTGPU_TRACE_CTAS_PER_SM"] = str(ctas)
            os.environ["RTGPU_LANES"] = str(lanes)
            os.environ["RTGPU_BATCH_MPATHS"] = str(batch)
            best = None
            for _ in range(args.reps + 1):  # first one is the warm-up
                accum.zero_()
                s0 = gpu.stats()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                gpu.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream)
                e1.record(stream)
                torch.cuda.synchronize()
                s1 = gpu.stats()
                rec = {"ms": e0.elapsed_time(e1), "trace": s1.ms_trace - s0.ms_trace, "shade": s1.ms_shade - s0.ms_shade, "raygen": s1.ms_raygen - s0.ms_raygen,
                       "rays": s1.rays - s0.rays, "traversed": s1.rays_traversed - s0.rays_traversed,
                       "trav_b": [int(s1.rays_traversed_per_bounce[i] - s0.rays_traversed_per_bounce[i]) for i in range(4)]}
                if best is None or rec["ms"] < best["ms"]:
                    best = rec
            img = accum[..., :3].double().cpu().numpy() / (T * k * k)
            if ref is None:
                ref = img
            diff = float(np.sqrt(np.mean((img - ref) ** 2)))
            print(json.dumps({"scene": args.scene, "variant": v, "ctas_per_sm": ctas, "lanes": lanes, "batch_mpaths": batch, "ms": round(best["ms"], 2),
                              "trace_ms": round(best["trace"], 2), "shade_ms": round(best["shade"], 2), "raygen_ms": round(best["raygen"], 2),
                              "mrays_s": round(best["rays"] / best["ms"] / 1e3, 1), "traversed": best["traversed"], "traversed_b0_3": best["trav_b"],
                              "mpaths_s": round(W * H * T * k * k / best["ms"] / 1e3, 1), "mean": round(float(img.mean()), 5),
                              "rmse_vs_first": round(diff, 6)}), flush=True)


if __name__ == "__main__":
    main()
