"""Generates tests/golden/*.npz from the CPU oracle (run once, here; fixtures are committed).

The reference ships no golden vectors and cannot be built in this environment (no Rust
toolchain), so these fixtures pin the ORACLE's behaviour over time (regression guard) and give
the GPU tests data that does not depend on rebuilding the oracle.   python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from rust_raytracer_b200 import api  # noqa: E402
import oracle  # noqa: E402
from util import primary_rays  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

for name in ("cornell", "light_test", "perlin"):
    scene, cam, prm = api.load([api.scene_path(name)])
    o = oracle.OracleScene(scene)
    rays, _ = primary_rays(oracle, cam, 512, seed=11)
    hits = o.trace_primary(rays)
    xi = np.random.default_rng(13).uniform(0.02, 0.98, (512, 12))
    b = o.bounce(rays, xi, 0.25)
    np.savez_compressed(os.path.join(HERE, f"{name}_first_hit.npz"), rays=rays, hits=hits, xi=xi, bounce=b)
    print(name, "hits", int((hits["object"] >= 0).sum()), "of 512")
