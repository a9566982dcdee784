"""Python host side of the rtgpu backend: a thin mirror of the reference's driver
(src/main.rs:25-88) over the C ABI in include/rtgpu.h / rtgpu_host.h.

    cfg   = Config.from_args(["scenes/cornell", "-s=100", "-t=8"])     # Config::from_args
    scene = LoadedScene(cfg)                                            # SceneLoader::load / built-ins
    cam, params = scene.camera()                                        # Camera::new
    gpu   = GpuScene(scene)                                             # upload (flatten + wide BVH)
    buf   = gpu.render(cam, params)                                     # Camera::render -> Buffer
    rgb   = tonemap_srgb8(buf)                                          # Writer::save maths

There is no CPU fallback: GpuScene raises RtgpuError when librtgpu.so or a CUDA device
is missing. The CPU oracle lives in oracle/ and is only reachable from tests and bench.py.
"""
import ctypes as C
import os

import numpy as np

from . import _abi as abi

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_PKG_DIR)
LIB_PATH = os.path.join(_PKG_DIR, "librtgpu.so")
ASSET_ROOT = os.path.join(REPO_ROOT, "assets") + os.sep  # holds scenes/resource/... for built-in scenes


class RtgpuError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"rtgpu error {code}: {message}")
        self.code = code


_lib = None


def lib():
    """Load librtgpu.so (built in-tree by __graft_entry__.build / make). Fails loudly."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtgpuError(abi.ERR_NO_DEVICE, f"{LIB_PATH} is missing: run `python __graft_entry__.py` (build) first")
        l = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        abi.bind(l, abi.HOST_SYMBOLS)
        abi.bind(l, abi.DEVICE_SYMBOLS)
        l.rtgpu_set_image_decoder(_DECODER, None)
        _lib = l
    return _lib


def _check(rc):
    if rc != 0:
        raise RtgpuError(rc, (lib().rtgpu_last_error() or b"").decode("utf-8", "replace"))


_libc = C.CDLL(None)
_libc.malloc.restype = C.c_void_p
_libc.malloc.argtypes = [C.c_size_t]


@abi.IMAGE_DECODE_FN
def _DECODER(path, out, _user):
    """JPEG & co: decode with PIL, hand the texels over as RGB8/RGB16 (buffer.rs:30-48)."""
    try:
        from PIL import Image
        im = Image.open(path.decode())
        if im.mode in ("I;16", "I;16B", "I"):
            arr = np.asarray(im).astype(np.uint16)
            arr = np.repeat(arr[:, :, None], 3, axis=2)
            fmt = abi.IMG_RGB16
        else:
            arr = np.asarray(im.convert("RGB"), dtype=np.uint8)
            fmt = abi.IMG_RGB8
        arr = np.ascontiguousarray(arr)
        mem = _libc.malloc(arr.nbytes)
        C.memmove(mem, arr.ctypes.data, arr.nbytes)
        out.contents.width, out.contents.height = arr.shape[1], arr.shape[0]
        out.contents.format = fmt
        out.contents.data = mem
        return 0
    except Exception:  # noqa: BLE001 - reported through the C error channel
        return -5


class Config:
    """Config::from_args (src/config.rs:62-176)."""

    def __init__(self, raw):
        self.raw = raw

    @classmethod
    def from_args(cls, args, program="rust_raytracer"):
        argv = [program.encode()] + [a.encode() for a in args]
        arr = (C.c_char_p * len(argv))(*argv)
        raw = abi.Config()
        _check(lib().rtgpu_config_from_args(len(argv), arr, C.byref(raw)))
        return cls(raw)

    @property
    def spp(self):
        return self.raw.thread_count * self.raw.sqrt_samples_per_thread ** 2


class LoadedScene:
    """main.rs:30-59: built-in scene names or a DSL file -> flat scene tables."""

    def __init__(self, cfg, asset_root=ASSET_ROOT):
        self.cfg = cfg
        self._h = C.c_void_p()
        _check(lib().rtgpu_scene_load(C.byref(cfg.raw), asset_root.encode(), C.byref(self._h)))
        self.desc = lib().rtgpu_loaded_scene_desc(self._h)
        self.warnings = (lib().rtgpu_loaded_scene_warnings(self._h) or b"").decode()

    def camera(self):
        cam, prm = abi.CameraDesc(), abi.RenderParams()
        _check(lib().rtgpu_camera_init(C.byref(self.cfg.raw), C.byref(cam), C.byref(prm)))
        return cam, prm

    def close(self):
        if self._h:
            lib().rtgpu_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def init_async(devices=None):
    """rtgpu_init_async: start CUDA bring-up on a helper thread; pair with GpuScene(scene, init=False)."""
    if devices is None:
        _check(lib().rtgpu_init_async(None, 0))
    else:
        arr = (C.c_int * len(devices))(*devices)
        _check(lib().rtgpu_init_async(arr, len(devices)))


def load(args, asset_root=ASSET_ROOT):
    """Convenience: argv-style flags -> (scene, camera, params)."""
    cfg = Config.from_args(args)
    scene = LoadedScene(cfg, asset_root)
    cam, prm = scene.camera()
    return scene, cam, prm


def scene_path(name):
    return os.path.join(REPO_ROOT, "assets", "scenes", name)


class GpuScene:
    """rtgpu_scene_create .. rtgpu_scene_destroy around an uploaded scene."""

    def __init__(self, scene, devices=None, init=True):
        self.scene = scene  # keep the tables alive
        if not init:
            pass  # init_async() is in flight: rtgpu_scene_create joins it after the host-side build
        elif devices is None:
            _check(lib().rtgpu_init(None, 0))
        else:
            arr = (C.c_int * len(devices))(*devices)
            _check(lib().rtgpu_init(arr, len(devices)))
        self._h = C.c_void_p()
        _check(lib().rtgpu_scene_create(scene.desc, C.byref(self._h)))

    def render(self, cam, params, out=None):
        """Camera::render: returns (H, W, 4) f64, the reference Buffer layout."""
        if out is None:
            out = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
        assert out.dtype == np.float64 and out.flags.c_contiguous
        _check(lib().rtgpu_render(self._h, C.byref(cam), C.byref(params), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def render_srgb8(self, cam, params, tonemap="aces"):
        """Camera::render + Writer::save maths, device-resident: returns (H, W, 3) u8."""
        out = np.zeros((cam.height, cam.width, 3), dtype=np.uint8)
        _check(lib().rtgpu_render_srgb8(self._h, C.byref(cam), C.byref(params), 1 if tonemap == "aces" else 0,
                                        out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def render_device(self, cam, params, accum_ptr, device_index=0, stream=None):
        _check(lib().rtgpu_render_device(self._h, device_index, C.byref(cam), C.byref(params),
                                         C.c_void_p(accum_ptr), C.c_void_p(stream or 0)))

    def trace_primary(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float64)
        n = rays.shape[0]
        hits = (abi.Hit * n)()
        _check(lib().rtgpu_trace_primary(self._h, rays.ctypes.data_as(C.POINTER(C.c_double)), n, hits))
        return hits_to_numpy(hits, n)

    def trace_fast(self, rays):
        """Closest hits from the production fp32 traversal kernel (parity hook)."""
        rays = np.ascontiguousarray(rays, dtype=np.float64)
        n = rays.shape[0]
        hits = (abi.FastHit * n)()
        _check(lib().rtgpu_trace_fast(self._h, rays.ctypes.data_as(C.POINTER(C.c_double)), n, hits))
        dt = np.dtype([("instance", np.int32), ("primitive", np.int32), ("t", np.float64), ("u", np.float32), ("v", np.float32)])
        return np.frombuffer(hits, dtype=dt, count=n).copy()

    def bounce(self, rays, xi, light_bias=0.25):
        """One bounce of ray_color with injected random numbers (parity hook)."""
        rays = np.ascontiguousarray(rays, dtype=np.float64)
        xi = np.ascontiguousarray(xi, dtype=np.float64)
        n = rays.shape[0]
        out = (abi.Bounce * n)()
        _check(lib().rtgpu_bounce(self._h, light_bias, rays.ctypes.data_as(C.POINTER(C.c_double)),
                                  xi.ctypes.data_as(C.POINTER(C.c_double)), xi.shape[1], n, out))
        return np.frombuffer(out, dtype=BOUNCE_DTYPE, count=n).copy()

    def stats(self):
        s = abi.Stats()
        _check(lib().rtgpu_get_stats(self._h, C.byref(s)))
        return s

    def close(self):
        if self._h:
            lib().rtgpu_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


HIT_DTYPE = np.dtype([("object", "<i4"), ("instance", "<i4"), ("primitive", "<i4"), ("front_face", "<i4"),
                      ("t", "<f8"), ("pos", "<f8", 3), ("normal", "<f8", 3), ("uv", "<f8", 2)])
assert HIT_DTYPE.itemsize == C.sizeof(abi.Hit)


BOUNCE_DTYPE = np.dtype([("kind", "<i4"), ("reserved", "<i4"), ("emission", "<f8", 3), ("weight", "<f8", 3),
                         ("next_origin", "<f8", 3), ("next_dir", "<f8", 3), ("pdf", "<f8"),
                         ("scattering_pdf", "<f8")])


def hits_to_numpy(hits, n):
    return np.frombuffer(hits, dtype=HIT_DTYPE, count=n).copy()


def resolve_device(accum_ptr, width, height, total_spp, device_index=0, stream=None, out=None):
    """K5: fp32 sums / spp ADDED into `out` (H, W, 4) f64 (a fresh zero buffer when None)."""
    if out is None:
        out = np.zeros((height, width, 4), dtype=np.float64)
    assert out.dtype == np.float64 and out.flags.c_contiguous
    _check(lib().rtgpu_resolve_device(device_index, C.c_void_p(accum_ptr), width, height, total_spp,
                                      out.ctypes.data_as(C.POINTER(C.c_double)), C.c_void_p(stream or 0)))
    return out


def generate_rays(cam, sqrt_spt, px, sxy, xi):
    """Camera::get_ray for injected random numbers (parity hook)."""
    px = np.ascontiguousarray(px, dtype=np.int32)
    sxy = np.ascontiguousarray(sxy, dtype=np.int32)
    xi = np.ascontiguousarray(xi, dtype=np.float64)
    n = px.shape[0]
    out = np.zeros((n, 6), dtype=np.float64)
    _check(lib().rtgpu_generate_rays(C.byref(cam), sqrt_spt, px.ctypes.data_as(C.POINTER(C.c_int32)),
                                     sxy.ctypes.data_as(C.POINTER(C.c_int32)),
                                     xi.ctypes.data_as(C.POINTER(C.c_double)), n,
                                     out.ctypes.data_as(C.POINTER(C.c_double))))
    return out


def tonemap_srgb8(buf, tonemap="aces"):
    """Writer::save maths (output.rs:23-50) on the GPU: (H, W, 4) f64 -> (H, W, 3) u8."""
    buf = np.ascontiguousarray(buf, dtype=np.float64)
    h, w = buf.shape[:2]
    out = np.zeros((h, w, 3), dtype=np.uint8)
    _check(lib().rtgpu_tonemap_srgb8(buf.ctypes.data_as(C.POINTER(C.c_double)), w, h,
                                     1 if tonemap == "aces" else 0, out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


def tonemap_srgb8_noalloc(buf, tonemap="aces"):
    """rtgpu_tonemap_srgb8_noalloc: the same maths through the library's static device scratch (no cudaMalloc);
    at most abi.NOALLOC_MAX_PIXELS pixels."""
    buf = np.ascontiguousarray(buf, dtype=np.float64)
    h, w = buf.shape[:2]
    out = np.zeros((h, w, 3), dtype=np.uint8)
    _check(lib().rtgpu_tonemap_srgb8_noalloc(buf.ctypes.data_as(C.POINTER(C.c_double)), w, h,
                                             1 if tonemap == "aces" else 0, out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


def save_png(path, rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    _check(lib().rtgpu_save_png(path.encode(), rgb.ctypes.data_as(C.POINTER(C.c_uint8)), rgb.shape[1], rgb.shape[0]))
