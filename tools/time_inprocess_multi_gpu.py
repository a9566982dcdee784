import sys, time, os
sys.path.insert(0, os.getcwd())
import numpy as np
from rust_raytracer_b200 import api
from tools.make_dragon_proxy import ensure
ensure()
n = int(sys.argv[1])
scene, cam, prm = api.load([api.scene_path("cornell_dragon"), "-w=1200", "-s=1000", "-t=10"])
gpu = api.GpuScene(scene, devices=list(range(n)))
prm.n_gpus = n
imgs = []
for i in range(4):
    t = time.perf_counter(); rgb = gpu.render_srgb8(cam, prm, "aces"); dt = time.perf_counter() - t
    imgs.append(rgb); print(f"in-process {n} GPU(s), render_srgb8 call {i}: {dt*1e3:.1f} ms", flush=True)
prm.n_gpus = 1
one = gpu.render_srgb8(cam, prm, "aces")
d = np.abs(one.astype(int) - imgs[-1].astype(int))
print("vs one GPU: max", d.max(), "differing", float((d > 0).mean()))
