"""TEST INFRASTRUCTURE — ctypes binding of the CPU oracle (oracle/liboracle.so).

The oracle is an f64 CPU restatement of the reference's path tracer. It is the checker
for tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs. Nothing under
rust_raytracer_b200/ imports this package: the product path has no CPU fallback.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from rust_raytracer_b200 import _abi as abi

_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_DIR, "liboracle.so")


class Bounce(C.Structure):
    _fields_ = [("kind", C.c_int32), ("reserved", C.c_int32), ("emission", C.c_double * 3),
                ("weight", C.c_double * 3), ("next_origin", C.c_double * 3), ("next_dir", C.c_double * 3),
                ("pdf", C.c_double), ("scattering_pdf", C.c_double)]


BOUNCE_DTYPE = np.dtype([("kind", "<i4"), ("reserved", "<i4"), ("emission", "<f8", 3), ("weight", "<f8", 3),
                         ("next_origin", "<f8", 3), ("next_dir", "<f8", 3), ("pdf", "<f8"),
                         ("scattering_pdf", "<f8")])

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _DIR])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        l = C.CDLL(LIB_PATH)
        dp = C.POINTER(C.c_double)
        l.oracle_last_error.restype = C.c_char_p
        l.oracle_scene_create.argtypes = [C.POINTER(abi.SceneDesc), C.POINTER(C.c_void_p)]
        l.oracle_scene_destroy.argtypes = [C.c_void_p]
        l.oracle_render.argtypes = [C.c_void_p, C.POINTER(abi.CameraDesc), C.POINTER(abi.RenderParams), C.c_int, dp,
                                    C.POINTER(C.c_uint64)]
        l.oracle_render_rows.argtypes = [C.c_void_p, C.POINTER(abi.CameraDesc), C.POINTER(abi.RenderParams), C.c_int,
                                         C.c_int, C.c_int, dp, C.POINTER(C.c_uint64)]
        l.oracle_trace_primary.argtypes = [C.c_void_p, dp, C.c_int64, C.POINTER(abi.Hit)]
        l.oracle_generate_rays.argtypes = [C.POINTER(abi.CameraDesc), C.c_int, C.POINTER(C.c_int32),
                                           C.POINTER(C.c_int32), dp, C.c_int64, dp]
        l.oracle_bounce.argtypes = [C.c_void_p, C.c_double, dp, dp, C.c_int, C.c_int64, C.POINTER(Bounce)]
        l.oracle_tonemap_srgb8.argtypes = [dp, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8)]
        l.oracle_aabb_test.argtypes = [dp, dp, dp, C.c_double, C.c_double]
        l.oracle_octree_stats.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        l.oracle_texture_sample.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, dp, dp]
        l.oracle_reflectance.argtypes = [C.c_double, C.c_double]
        l.oracle_reflectance.restype = C.c_double
        _lib = l
    return _lib


class OracleError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise OracleError((lib().oracle_last_error() or b"").decode())


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class OracleScene:
    def __init__(self, loaded_scene):
        self.loaded = loaded_scene
        self._h = C.c_void_p()
        _check(lib().oracle_scene_create(loaded_scene.desc, C.byref(self._h)))

    def render(self, cam, params, n_threads=0, out=None):
        """Camera::render -> ((H, W, 4) f64 buffer, world.test call count)."""
        if out is None:
            out = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
        rays = C.c_uint64(0)
        _check(lib().oracle_render(self._h, C.byref(cam), C.byref(params), n_threads, _dp(out), C.byref(rays)))
        return out, rays.value

    def render_rows(self, cam, params, row_offset, row_step, n_threads=0):
        """Bounded sample of Camera::render: only rows row_offset (mod row_step)."""
        out = np.zeros((cam.height, cam.width, 4), dtype=np.float64)
        rays = C.c_uint64(0)
        _check(lib().oracle_render_rows(self._h, C.byref(cam), C.byref(params), n_threads, row_offset, row_step,
                                        _dp(out), C.byref(rays)))
        return out, rays.value

    def trace_primary(self, rays):
        from rust_raytracer_b200.api import hits_to_numpy
        rays = np.ascontiguousarray(rays, dtype=np.float64)
        n = rays.shape[0]
        hits = (abi.Hit * n)()
        _check(lib().oracle_trace_primary(self._h, _dp(rays), n, hits))
        return hits_to_numpy(hits, n)

    def bounce(self, rays, xi, light_bias=0.25):
        rays = np.ascontiguousarray(rays, dtype=np.float64)
        xi = np.ascontiguousarray(xi, dtype=np.float64)
        n = rays.shape[0]
        out = (Bounce * n)()
        _check(lib().oracle_bounce(self._h, light_bias, _dp(rays), _dp(xi), xi.shape[1], n, out))
        return np.frombuffer(out, dtype=BOUNCE_DTYPE, count=n).copy()

    def octree_stats(self, mesh_index=0):
        out = (C.c_int64 * 4)()
        _check(lib().oracle_octree_stats(self._h, mesh_index, out))
        return dict(nodes=out[0], leaves=out[1], leaf_entries=out[2], max_depth=out[3])

    def texture_sample(self, tex, u, v, p):
        p = np.asarray(p, dtype=np.float64)
        out = np.zeros(3)
        _check(lib().oracle_texture_sample(self._h, tex, u, v, _dp(p), _dp(out)))
        return out

    def close(self):
        if self._h:
            lib().oracle_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def generate_rays(cam, sqrt_spt, px, sxy, xi):
    px = np.ascontiguousarray(px, dtype=np.int32)
    sxy = np.ascontiguousarray(sxy, dtype=np.int32)
    xi = np.ascontiguousarray(xi, dtype=np.float64)
    n = px.shape[0]
    out = np.zeros((n, 6), dtype=np.float64)
    _check(lib().oracle_generate_rays(C.byref(cam), sqrt_spt, px.ctypes.data_as(C.POINTER(C.c_int32)),
                                      sxy.ctypes.data_as(C.POINTER(C.c_int32)), _dp(xi), n, _dp(out)))
    return out


def tonemap_srgb8(buf, tonemap="aces"):
    buf = np.ascontiguousarray(buf, dtype=np.float64)
    h, w = buf.shape[:2]
    out = np.zeros((h, w, 3), dtype=np.uint8)
    _check(lib().oracle_tonemap_srgb8(_dp(buf), w, h, 1 if tonemap == "aces" else 0,
                                      out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


def aabb_test(lo, hi, ray, t_lo, t_hi):
    lo, hi, ray = (np.asarray(a, dtype=np.float64) for a in (lo, hi, ray))
    return bool(lib().oracle_aabb_test(_dp(lo), _dp(hi), _dp(ray), t_lo, t_hi))


def reflectance(cos_theta, ior_ratio):
    return lib().oracle_reflectance(cos_theta, ior_ratio)
