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


def load_coverage_scene(api, extra=()):
    """assets/scenes/coverage_test (this repository's own scene: constant media behind sphere boundaries, isotropic,
    normal_debug, checker_solid, uv_debug, a bare mesh) plus what the DSL cannot express, patched into the C-ABI
    tables the way a host would set it: Clamp mode on the albedo image (image.rs:44), uvs of the second mesh
    stretched to [-0.5, 1.5] so that Clamp and Repeat differ, flat_shading + hit_back_faces on it (mesh.rs:25-26)."""
    from rust_raytracer_b200 import _abi as abi
    scene, cam, prm = api.load([api.scene_path("coverage_test")] + list(extra))
    d = scene.desc.contents
    imgs = [i for i in range(d.n_textures) if d.textures[i].kind == abi.TEX_IMAGE]
    assert len(imgs) == 2
    d.textures[imgs[1]].repeat = 0                       # mat_img (declared after mat_dbg_map)
    meshes = [i for i in range(d.n_objects) if d.objects[i].kind == abi.OBJ_MESH]
    assert len(meshes) == 2
    d.objects[meshes[1]].flags |= abi.OBJF_FLAT_SHADING | abi.OBJF_HIT_BACK_FACES
    m = d.meshes[d.objects[meshes[1]].mesh]
    assert d.objects[meshes[0]].mesh != d.objects[meshes[1]].mesh
    uv = np.ctypeslib.as_array(m.uvs, shape=(m.n_uvs, 2))
    if uv.max() <= 1.0 + 1e-9:                           # (idempotent: the loader's table is patched once)
        uv[:] = uv * 2.0 - 0.5
    return scene, cam, prm
