#!/usr/bin/env python
"""tools/time_build.py — host-side "Ready" phase of a scene (no GPU needed): scene DSL + OBJ parse, then
flatten + wide-BVH build through the host-simulation library (same flatten.cpp / bvh.cpp as librtgpu.so).
RTGPU_BUILD_TIMING=1 prints the cumulative time after every stage."""
import ctypes as C
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from rust_raytracer_b200 import api
    from tools.make_dragon_proxy import ensure
    ensure()
    name = sys.argv[1] if len(sys.argv) > 1 else "cornell_dragon"
    out = os.path.join(ROOT, "build", "libhostsim.so")
    dev = os.path.join(ROOT, "rust_raytracer_b200", "csrc", "device")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-I" + os.path.join(ROOT, "include"),
                           "-I" + dev, "-x", "c++", os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp"),
                           os.path.join(dev, "flatten.cpp"), os.path.join(dev, "bvh.cpp"), "-o", out])
    sim = C.CDLL(out)
    os.environ["RTGPU_BUILD_TIMING"] = "1"
    for rep in range(2):
        t0 = time.perf_counter()
        scene, cam, prm = api.load([api.scene_path(name)])
        t1 = time.perf_counter()
        h = C.c_void_p()
        assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0
        t2 = time.perf_counter()
        print(f"rep {rep}: load (DSL + OBJ) {1e3 * (t1 - t0):.0f} ms, flatten + BVH {1e3 * (t2 - t1):.0f} ms", flush=True)
        sim.hostsim_scene_destroy(h)


if __name__ == "__main__":
    main()
