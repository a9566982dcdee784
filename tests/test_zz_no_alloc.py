"""rtgpu_tonemap_srgb8_noalloc — the entry point for callers that may not allocate (no counterpart in the reference).
Collected last on purpose.

The static device scratch is shared by every call and keeps the last call's contents, so the check is two calls in a
row with DIFFERENT images (the second smaller: stale bytes of the first would show), each against the existing
rtgpu_tonemap_srgb8, within the tolerance of test_tonemap_matches_oracle (one 8-bit step on < 0.1 % of the values; the
two entry points run the same kernel, so in practice the outputs are identical)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _image(rng, h, w):
    buf = np.zeros((h, w, 4))
    buf[..., :3] = rng.gamma(0.5, 2.0, (h, w, 3))
    buf[0, 0, :3] = [0, 1e-4, 1e6]
    return buf


def test_no_alloc_tonemap_matches_the_allocating_entry_point_twice_in_a_row(api):
    from rust_raytracer_b200 import _abi as abi
    rng = np.random.default_rng(2)
    for h, w, mode in ((512, 512, "aces"), (97, 61, "clamp"), (64, 64, "aces")):   # first call: exactly the maximum size
        assert h * w <= abi.NOALLOC_MAX_PIXELS
        buf = _image(rng, h, w)
        want = api.tonemap_srgb8(buf, mode)
        got = api.tonemap_srgb8_noalloc(buf, mode)
        assert got.shape == want.shape == (h, w, 3) and got.max() > 100
        d = np.abs(got.astype(int) - want.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3, (h, w, mode, d.max(), (d > 0).mean())


def test_no_alloc_tonemap_rejects_images_beyond_the_static_scratch(api):
    buf = np.zeros((513, 512, 4))
    with pytest.raises(api.RtgpuError) as e:
        api.tonemap_srgb8_noalloc(buf)
    assert e.value.code == -1 and "RTGPU_NOALLOC_MAX_PIXELS" in str(e.value)
