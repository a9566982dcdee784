"""Built-in image decoders of the host library (csrc/host/png.cpp, jpeg.cpp) behind rtgpu_decode_image: what an
`image <file>` texture holds (texture/image.rs:21-33, buffer.rs:30-48) before into_rgb32f's /255.

The reference decodes through the `image` crate (third-party, absent, parity unpinned): the built-in decoders are checked
against PIL (libpng / libjpeg-turbo) on files written here — PNG bit-exact, JPEG to the +-1 of two correct inverse DCTs
(a few levels after the colour transform) — and on the one JPEG of the BASELINE scenes, earthmap.jpg."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PIL_Image = pytest.importorskip("PIL.Image")

_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]


def _decode(api, path, builtin_only=1):
    from rust_raytracer_b200 import _abi as abi
    im = abi.Image()
    rc = api.lib().rtgpu_decode_image(str(path).encode(), builtin_only, C.byref(im))
    if rc != 0:
        return rc, (api.lib().rtgpu_last_error() or b"").decode()
    dt, ct = (np.uint8, C.c_uint8) if im.format == abi.IMG_RGB8 else (np.uint16, C.c_uint16)
    n = im.width * im.height * 3
    a = np.ctypeslib.as_array(C.cast(im.data, C.POINTER(ct)), shape=(n,)).reshape(im.height, im.width, 3).astype(dt)
    _libc.free(im.data)
    return 0, a


def _picture(h=203, w=317, seed=0):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([128 + 100 * np.sin(xx / 17.0) * np.cos(yy / 23.0), 128 + 90 * np.cos(xx / 9.0 + yy / 31.0), (xx * yy) % 256], -1)
    return np.clip(img + rng.normal(0, 6, img.shape), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("name,kw", [
    ("444", dict(subsampling=0)), ("422", dict(subsampling=1)), ("420", dict(subsampling=2)),
    ("420_q35", dict(subsampling=2, quality=35)), ("444_optimised_tables", dict(subsampling=0, optimize=True)),
    ("grey", dict()), ("restart_rows", dict(subsampling=2, restart_marker_rows=1)), ("restart_blocks", dict(subsampling=0, restart_marker_blocks=5)),
    ("tiny_7x5", dict(subsampling=2)), ("one_pixel", dict(subsampling=0)),
])
def test_builtin_jpeg_decoder_agrees_with_libjpeg(api, tmp_path, name, kw):
    kw = dict(kw)
    img = _picture()
    if name == "tiny_7x5":
        img = img[:5, :7]
    if name == "one_pixel":
        img = img[:1, :1]
    p = tmp_path / (name + ".jpg")
    PIL_Image.fromarray(img[..., 0] if name == "grey" else img).save(p, quality=kw.pop("quality", 90), **kw)
    rc, got = _decode(api, p)
    assert rc == 0, got
    ref = np.asarray(PIL_Image.open(p).convert("RGB"))
    assert got.shape == ref.shape and got.dtype == np.uint8
    d = np.abs(got.astype(int) - ref.astype(int))
    # two correct IDCTs differ by one level in a few samples; the colour transform can stretch that to 3
    assert d.max() <= 4 and d.mean() <= 0.15, (name, d.max(), d.mean())


def test_builtin_jpeg_decoder_on_the_baseline_texture(api):
    """scenes/earth's earthmap.jpg (C3a): SOF0, 4:4:4, 1024x512."""
    p = os.path.join(ROOT, "assets", "scenes", "resource", "earthmap.jpg")
    rc, got = _decode(api, p)
    assert rc == 0, got
    ref = np.asarray(PIL_Image.open(p).convert("RGB"))
    d = np.abs(got.astype(int) - ref.astype(int))
    assert got.shape == (512, 1024, 3) and d.max() <= 4 and d.mean() <= 0.05
    # with the plug-in (PIL, registered by the Python front-end) the same call returns PIL's texels exactly
    rc, via_plugin = _decode(api, p, builtin_only=0)
    assert rc == 0 and np.array_equal(via_plugin, ref)


def test_builtin_jpeg_decoder_refuses_what_it_does_not_implement(api, tmp_path):
    p = tmp_path / "progressive.jpg"
    PIL_Image.fromarray(_picture()).save(p, progressive=True)
    rc, msg = _decode(api, p)
    assert rc != 0 and "progressive" in msg
    q = tmp_path / "truncated.jpg"
    PIL_Image.fromarray(_picture()).save(q)
    data = q.read_bytes()
    q.write_bytes(data[:200])
    rc, msg = _decode(api, q)
    assert rc != 0 and "jpeg" in msg
    rc, msg = _decode(api, tmp_path / "missing.jpg")
    assert rc != 0 and "cannot open" in msg
    r = tmp_path / "not_a.jpg"
    r.write_bytes(b"hello, this is not a JPEG file at all")
    rc, msg = _decode(api, r)
    assert rc != 0 and "not a JPEG" in msg


@pytest.mark.parametrize("mode", ["RGB", "L", "RGBA", "LA", "P", "I;16"])
def test_builtin_png_decoder_is_exact(api, tmp_path, mode):
    img = _picture(61, 45)
    p = tmp_path / "t.png"
    if mode == "RGB":
        PIL_Image.fromarray(img).save(p)
        ref = img
    elif mode == "L":
        PIL_Image.fromarray(img[..., 0]).save(p)
        ref = np.repeat(img[..., :1], 3, axis=2)
    elif mode == "RGBA":
        PIL_Image.fromarray(np.dstack([img, img[..., 1]])).save(p)
        ref = img  # into_rgb32f drops alpha
    elif mode == "LA":
        PIL_Image.fromarray(np.dstack([img[..., 0], img[..., 1]]), "LA").save(p)
        ref = np.repeat(img[..., :1], 3, axis=2)
    elif mode == "P":
        pal = PIL_Image.fromarray(img).quantize(64)
        pal.save(p)
        ref = np.asarray(pal.convert("RGB"))
    else:
        g16 = (img[..., 0].astype(np.uint16) << 8) | img[..., 1]
        PIL_Image.fromarray(g16).save(p)
        ref = np.repeat(g16[..., None], 3, axis=2)
    rc, got = _decode(api, p)
    assert rc == 0, got
    assert got.shape == ref.shape and np.array_equal(got, ref)


def test_scene_with_a_jpeg_texture_loads_without_a_plugin(api):
    """scenes/earth through the library alone (what the native CLI does): the plug-in unregistered, the texture table of the
    loaded scene holds the built-in decoder's texels."""
    from rust_raytracer_b200 import _abi as abi
    lib = api.lib()
    lib.rtgpu_set_image_decoder(C.cast(None, abi.IMAGE_DECODE_FN), None)
    try:
        scene, cam, prm = api.load([api.scene_path("earth"), "-w=64", "-s=4", "-t=1"])
        d = scene.desc.contents
        assert d.n_images == 1
        im = d.images[0]
        assert (im.width, im.height, im.format) == (1024, 512, abi.IMG_RGB8)
        texels = np.ctypeslib.as_array(C.cast(im.data, C.POINTER(C.c_uint8)), shape=(512 * 1024 * 3,)).reshape(512, 1024, 3)
        ref = np.asarray(PIL_Image.open(os.path.join(ROOT, "assets", "scenes", "resource", "earthmap.jpg")).convert("RGB"))
        assert np.abs(texels.astype(int) - ref.astype(int)).max() <= 4
        scene.close()
    finally:
        lib.rtgpu_set_image_decoder(api._DECODER, None)
