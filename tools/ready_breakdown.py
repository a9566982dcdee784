import time, os, sys
sys.path.insert(0, "/root/repo")
t0 = time.perf_counter()
from rust_raytracer_b200 import api
import ctypes as C
t1 = time.perf_counter()
api.lib()
t2 = time.perf_counter()
os.environ["RTGPU_BUILD_TIMING"] = "1"
arr = (C.c_int * 1)(0)
api._check(api.lib().rtgpu_init(arr, 1))
t3 = time.perf_counter()
scene, cam, prm = api.load([api.scene_path("cornell_dragon"), "-w=1200", "-s=1000", "-t=10"])
t4 = time.perf_counter()
g = api.GpuScene(scene, devices=[0])
t5 = time.perf_counter()
import numpy as np
prm.sample_end = 10
buf = g.render(cam, prm)
t6 = time.perf_counter()
buf = g.render(cam, prm)
t7 = time.perf_counter()
print(f"import {t1-t0:.3f} lib {t2-t1:.3f} init {t3-t2:.3f} load {t4-t3:.3f} scene_create {t5-t4:.3f} first render(10spp) {t6-t5:.3f} second {t7-t6:.3f}")
