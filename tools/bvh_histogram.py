#!/usr/bin/env python
"""tools/bvh_histogram.py [scene] — fill statistics of the 8-wide BVH (host build, no GPU needed).

Builds tests/hostsim (the host-compiled copy of flatten.cpp + bvh.cpp) and prints how many slots the
nodes use, how many primitives the leaf slots hold and the build time.
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from rust_raytracer_b200 import api
    from tools.make_dragon_proxy import ensure
    ensure()
    name = sys.argv[1] if len(sys.argv) > 1 else "cornell_dragon"
    out = os.path.join(ROOT, "build", "libhostsim.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    dev = os.path.join(ROOT, "rust_raytracer_b200", "csrc", "device")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-I" + os.path.join(ROOT, "include"),
                           "-I" + dev, "-x", "c++", os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp"),
                           os.path.join(dev, "flatten.cpp"), os.path.join(dev, "bvh.cpp"), "-o", out])
    sim = C.CDLL(out)
    scene, cam, prm = api.load([] if name == "default" else [api.scene_path(name)])
    h = C.c_void_p()
    assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0
    info, hist = (C.c_int64 * 4)(), (C.c_int64 * 16)()
    sim.hostsim_bvh_info(h, info)
    sim.hostsim_bvh_histogram(h, hist)
    nodes, prims = info[0], info[1]
    print(f"{name}: {nodes} nodes, {prims} prims, depth {info[2]}, build {info[3]} ms, {prims / max(nodes, 1):.2f} prims/node")
    print("slots used per node:", {k: hist[k] for k in range(9)}, "mean", sum(k * hist[k] for k in range(9)) / max(nodes, 1))
    print("leaf slots with 1/2/3 prims:", hist[10], hist[11], hist[12], " internal child slots:", hist[13])


if __name__ == "__main__":
    main()
