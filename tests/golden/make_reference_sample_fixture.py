"""Generates tests/golden/reference_sample1_blocks.npz from the REFERENCE's own output image.

samples/sample1.png (README.md:12) is what the reference binary wrote for BASELINE config 4,
`rust_raytracer scenes/light_test -w=2400 -s=1000 -t=10` (2400x1600 @1000spp, ~20 min on an M3 Pro): a fully
deterministic scene (no random geometry), so — unlike anything the reference's RNG touches — its converged image
is reproducible up to Monte-Carlo noise. The PNG itself (5.6 MB) stays in /root/reference; this script (run once,
in the build container) commits its 8x8 and 10x10 block means (sRGB values, 0..1, stored as uint16 / 65535) and
its channel means. They are the one place where this repository is pinned against an OUTPUT OF THE REFERENCE
BINARY rather than against the restated oracle.   python tests/golden/make_reference_sample_fixture.py
"""
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/samples/sample1.png"

im = np.asarray(Image.open(SRC).convert("RGB")).astype(np.float64) / 255.0
h, w = im.shape[:2]
assert (w, h) == (2400, 1600)
out = {"channel_mean": im.mean(axis=(0, 1))}
for b in (8, 10):
    blocks = im.reshape(h // b, b, w // b, b, 3).mean(axis=(1, 3))
    out[f"blocks{b}"] = np.round(blocks * 65535.0).astype(np.uint16)
np.savez_compressed(os.path.join(HERE, "reference_sample1_blocks.npz"), **out)
print("channel mean", out["channel_mean"], {k: v.shape for k, v in out.items()})
