#!/usr/bin/env python
"""tools/sort_experiment.py — how much would ray reordering buy k_trace_fast? (GPU box only)

Builds realistic bounce-1..3 ray sets of cornell_dragon with the parity hooks (rtgpu_bounce with
uniform draws), then times the production trace kernel alone (RTGPU_TIME_HOOK=1) on the same rays in
different queue orders: as generated (pixel order), random permutation, Morton order of the origin,
Morton of origin + direction octant, and sorted by the primitive the ray starts on (prim id of the
previous hit, BVH leaf order). Only rays that reach the BVH bounds are kept (the others bypass the
kernel in a render).
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def morton3(q):
    """q: (n, 3) uint32 with 10 bits each -> 30-bit Morton code."""
    def spread(x):
        x = x.astype(np.uint64) & 0x3FF
        x = (x | (x << 16)) & 0x30000FF
        x = (x | (x << 8)) & 0x300F00F
        x = (x | (x << 4)) & 0x30C30C3
        x = (x | (x << 2)) & 0x9249249
        return x
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=3_000_000)
    ap.add_argument("--scene", default="cornell_dragon")
    ap.add_argument("--width", type=int, default=1200)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    os.environ["RTGPU_TIME_HOOK"] = "1"
    from rust_raytracer_b200 import api
    from tools.make_dragon_proxy import ensure
    ensure()
    scene, cam, prm = api.load([api.scene_path(args.scene), f"-w={args.width}", "-t=10", "-s=10", "--max-depth=20", "--light-bias=0.25"])
    gpu = api.GpuScene(scene, devices=[0])
    rng = np.random.default_rng(5)
    n = args.n
    # primary rays in scanline order (what k_raygen emits)
    idx = np.sort(rng.integers(0, cam.width * cam.height, n))
    px = np.stack([idx % cam.width, idx // cam.width], 1).astype(np.int32)
    rays = api.generate_rays(cam, 1, px, np.zeros((n, 2), np.int32), rng.random((n, 4)))

    def time_order(rr, order, label, bounce):
        r = rr[order]
        best = 1e9
        for _ in range(args.reps):
            h = gpu.trace_fast(r)
            best = min(best, gpu.stats().ms_trace)
        print(json.dumps({"bounce": bounce, "order": label, "rays": len(r), "trace_ms": round(best, 3),
                          "mrays_s": round(len(r) / best / 1e3, 1)}), flush=True)
        return h

    for bounce in range(args.bounces + 1):
        rays[:, 3:6] = rays[:, 3:6].astype(np.float32)
        m = len(rays)
        # keep the rays that walk the BVH: instance >= 0 of the mesh is not known before tracing, so trace once
        h = gpu.trace_fast(rays)
        if bounce == 0:
            # mesh bounds from the hit points of the instance with the most distinct primitives
            insts = [i for i in np.unique(h["instance"]) if i >= 0]
            mesh = max(insts, key=lambda i: len(np.unique(h["primitive"][h["instance"] == i])))
            sel = h["instance"] == mesh
            pts = rays[sel, 0:3] + rays[sel, 3:6] * h["t"][sel, None]
            blo, bhi = pts.min(0), pts.max(0)
            print(json.dumps({"mesh_instance": int(mesh), "bounds_lo": blo.tolist(), "bounds_hi": bhi.tolist()}), flush=True)
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1.0 / rays[:, 3:6]
            t0 = (blo - rays[:, 0:3]) * inv; t1 = (bhi - rays[:, 0:3]) * inv
            tn = np.nanmax(np.minimum(t0, t1), axis=1); tf = np.nanmin(np.maximum(t0, t1), axis=1)
        reach = (tn <= tf) & (tf > 0.001) & (tn < np.where(np.isfinite(h["t"]) & (h["instance"] >= 0), h["t"], np.inf))
        all_rays = rays
        rays = rays[reach]
        m = len(rays)
        lo = rays[:, 0:3].min(0); hi = rays[:, 0:3].max(0)
        q = np.clip((rays[:, 0:3] - lo) / np.maximum(hi - lo, 1e-9) * 1023.0, 0, 1023).astype(np.uint32)
        mo = morton3(q)
        octant = ((rays[:, 3] < 0).astype(np.uint64) << 2) | ((rays[:, 4] < 0).astype(np.uint64) << 1) | (rays[:, 5] < 0).astype(np.uint64)
        ident = np.arange(m)
        time_order(rays, ident, "as generated", bounce)
        time_order(rays, rng.permutation(m), "random permutation", bounce)
        time_order(rays, np.argsort(mo, kind="stable"), "morton(origin) 30 bit", bounce)
        time_order(rays, np.argsort(mo >> 15, kind="stable"), "morton(origin) 15 bit", bounce)
        time_order(rays, np.argsort((mo >> 9) << 3 | octant, kind="stable"), "morton(origin) 21 bit + octant", bounce)
        time_order(rays, np.argsort(octant << 30 | mo, kind="stable"), "octant major, morton(origin) minor", bounce)
        if bounce == args.bounces:
            break
        # next bounce through the scatter hook; ray order keeps the "as generated" (retire-like) order
        xi = rng.uniform(0.0, 1.0, (len(all_rays), 12))
        b = gpu.bounce(all_rays, xi, 0.25)
        cont = np.isin(b["kind"], [1, 2]) & np.isfinite(b["next_dir"]).all(axis=1) & (np.linalg.norm(b["next_dir"], axis=1) > 1e-9)
        rays = np.concatenate([b["next_origin"][cont], b["next_dir"][cont]], axis=1)
        # scramble mildly like the pipeline does (queue append order = retirement order): shuffle inside windows of 4096
        w = 4096
        perm = np.arange(len(rays))
        for s in range(0, len(rays), w):
            rng.shuffle(perm[s:s + w])
        rays = rays[perm]


if __name__ == "__main__":
    main()
