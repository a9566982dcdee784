"""Shared helpers for the parity tests."""
import numpy as np


def primary_rays(oracle, cam, n, seed=0, sqrt_spt=1):
    """n camera rays through random pixels with random jitter / lens draws (camera.rs:260-280)."""
    rng = np.random.default_rng(seed)
    px = np.stack([rng.integers(0, cam.width, n), rng.integers(0, cam.height, n)], 1).astype(np.int32)
    sxy = rng.integers(0, sqrt_spt, (n, 2)).astype(np.int32)
    xi = np.concatenate([rng.random((n, 2)), rng.standard_normal((n, 2))], 1)
    return oracle.generate_rays(cam, sqrt_spt, px, sxy, xi), (px, sxy, xi)


def luminance(buf):
    return 0.2126 * buf[..., 0] + 0.7152 * buf[..., 1] + 0.0722 * buf[..., 2]


def rmse(a, b):
    return float(np.sqrt(np.mean((a[..., :3] - b[..., :3]) ** 2)))
