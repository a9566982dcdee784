"""Generate the deterministic stand-in for the missing `resource/dragon_high.obj`
(.MISSING_LARGE_BLOBS of the reference): monkey.obj subdivided 1->4 three times
(15 744 -> 1 007 616 triangles) with a smooth procedural displacement, rescaled to 5.8 object units wide and resting on
y = 0 like the dragon in samples/sample2.png. PROXY MESH."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def ensure(levels=3, displacement=0.004, target_width=5.8):
    from rust_raytracer_b200 import api
    res = os.path.join(ROOT, "assets", "scenes", "resource")
    dst = os.path.join(res, "dragon_high.obj")
    if not os.path.exists(dst):
        api._check(api.lib().rtgpu_make_proxy_mesh(os.path.join(res, "monkey.obj").encode(), dst.encode(), levels, displacement, target_width))
    return dst


if __name__ == "__main__":
    path = ensure()
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    print(path, os.path.getsize(path), "bytes sha256", h.hexdigest())
