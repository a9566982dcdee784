"""rust_raytracer_b200 — B200-native path-tracing backend for teofum/rust_raytracer.

Only what the hot path (Camera::render) needs: the C-ABI library (csrc/, librtgpu.so)
and this thin Python mirror of the reference's driver.
"""
from . import _abi as abi  # noqa: F401
