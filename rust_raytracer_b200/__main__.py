"""`python -m rust_raytracer_b200 [<scene>] [-w= -r= -s= -a= -f= -t= ... --max-depth= --light-bias=] --gpu`

The reference's driver (src/main.rs:25-88) with the one call site `camera.render(world, lights, &mut buf)`
(main.rs:74-75) replaced by rtgpu_render: same positional scene argument (built-in name or DSL file), same
`-k=v` flags and defaults (config.rs:62-176, parsed by the C++ host library), same three progress lines, same
output file `out.png` written through the Writer::save maths (ACES -> sRGB OETF -> trunc(x * 255.999),
output.rs:23-50) on the GPU.

Flags the reference does not have (all `--name[=value]`, removed before Config::from_args sees the rest):
  --gpu        accepted for symmetry with the north-star command line; this backend is GPU-only and fails loudly
               when the CUDA library or a device is missing (there is no CPU fallback)
  --gpus=N     shard the sample strata over N devices of this process (rtgpu_render, peer-read reduce)
  --seed=N     Philox key of the sampling streams (the reference seeds from OS entropy, camera.rs:208-209)
  --out=PATH   output file (default out.png like main.rs:23)
"""
import ctypes as C
import sys
import time

import numpy as np


def _ensure_dragon_proxy(api, args):
    """scenes/cornell_dragon names resource/dragon_high.obj, a large blob missing from the reference tree
    (.MISSING_LARGE_BLOBS): generate the deterministic ~1.0 M-triangle stand-in next to the scene file once
    (same generator and parameters as tools/make_dragon_proxy.py) and say so."""
    import os
    for a in args:
        if a.startswith("-") or not os.path.isfile(a):
            continue
        with open(a, "r", errors="replace") as f:
            if "resource/dragon_high.obj" not in f.read():
                continue
        res = os.path.join(os.path.dirname(os.path.abspath(a)), "resource")
        dst, src = os.path.join(res, "dragon_high.obj"), os.path.join(res, "monkey.obj")
        if not os.path.exists(dst) and os.path.exists(src):
            print(f"note: {dst} is missing upstream; generating the proxy mesh (subdivided, displaced monkey.obj)", file=sys.stderr)
            api._check(api.lib().rtgpu_make_proxy_mesh(src.encode(), dst.encode(), 3, C.c_double(0.004), C.c_double(5.8)))


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    n_gpus, seed, out_path, preview = 1, 1, "out.png", 0
    rest = []
    for a in argv:
        if a == "--gpu":
            continue
        if a.startswith("--gpus="):
            n_gpus = int(a.split("=", 1)[1])
        elif a.startswith("--seed="):
            seed = int(a.split("=", 1)[1])
        elif a.startswith("--out="):
            out_path = a.split("=", 1)[1]
        elif a.startswith("--preview="):      # progressive output: the PNG is rewritten after every N samples per pixel
            preview = max(0, int(a.split("=", 1)[1]))
        else:
            rest.append(a)
    from . import api

    _ensure_dragon_proxy(api, rest)
    t0 = time.perf_counter()
    api.init_async(list(range(n_gpus)))                   # CUDA bring-up on a helper thread, overlapped with everything below
    try:
        scene, cam, prm = api.load(rest)                  # main.rs:26-59 (Config::from_args + scene construction)
    except api.RtgpuError as e:
        raise SystemExit(f"rust_raytracer_b200: {e}")
    if scene.warnings:
        print(scene.warnings, file=sys.stderr)
    try:
        gpu = api.GpuScene(scene, init=False)             # flatten + wide BVH (host), join the bring-up, upload: part of "Ready"
    except api.RtgpuError as e:
        raise SystemExit(f"rust_raytracer_b200: {n_gpus} CUDA device(s) requested: {e}")
    print(f"Ready: {time.perf_counter() - t0:.2f}s")
    spp = prm.replicas * prm.sqrt_spt * prm.sqrt_spt
    print(f"Rendering: {cam.width}x{cam.height} @{spp}spp on {prm.replicas} threads ({prm.sqrt_spt * prm.sqrt_spt} samples/thread)"
          f" [{n_gpus} GPU(s)]")
    prm.n_gpus, prm.seed = n_gpus, seed
    # camera.create_buffer + camera.render + Writer::save's tonemap_aces / sRGB / u8 maths (main.rs:74-82) in one
    # device-resident call: 3 bytes per pixel come back instead of the 32-byte f64 Buffer going down and up again
    if preview > 0 and preview < spp:
        # Progressive: Camera::render ADDS each block's share (sum of its samples / spp, camera.rs:229,247-253) into the same
        # f64 Buffer; after m of spp samples the preview is that buffer scaled by spp / m. Same samples, same Philox streams
        # as the single call: the final image differs from it by fp32 / f64 summation order only.
        buf = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
        for begin in range(0, spp, preview):
            end = min(spp, begin + preview)
            blk = type(prm).from_buffer_copy(prm)
            blk.sample_begin, blk.sample_end = begin, end
            gpu.render(cam, blk, out=buf)
            rgb = api.tonemap_srgb8(buf * (spp / end), "aces")
            api.save_png(out_path, rgb)
            print(f"  {end}/{spp} spp: {time.perf_counter() - t0:.2f}s")
    else:
        rgb = gpu.render_srgb8(cam, prm, "aces")
    print(f"Done: {time.perf_counter() - t0:.2f}s. Writing output to file...")
    api.save_png(out_path, rgb)
    st = gpu.stats()
    print(f"Done! Took {time.perf_counter() - t0:.2f}s. Goodbye :)  "
          f"[{st.paths / 1e6:.1f} Mpaths, {st.rays / 1e6:.1f} Mrays, {st.kernel_launches} kernel launches]")
    return 0


if __name__ == "__main__":
    sys.exit(main())
