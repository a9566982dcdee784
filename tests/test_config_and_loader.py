"""Host front-end (CPU): Config::from_args, scene DSL, OBJ loader, Camera::init — semantics of
reference src/config.rs, src/loaders/scene.rs, src/loaders/obj.rs, src/camera.rs:47-130."""
import ctypes as C
import math
import os

import numpy as np
import pytest


def test_readme_spp_rule(api):
    # README.md:43-49: 10 threads at 100 spp -> 9 samples/thread (3x3) -> 90 spp
    cfg = api.Config.from_args(["-t=10", "-s=100"])
    assert cfg.raw.sqrt_samples_per_thread == 3 and cfg.spp == 90
    assert api.Config.from_args(["-t=1", "-s=1000"]).spp == 961      # 31^2
    assert api.Config.from_args(["-t=8", "-s=100"]).spp == 72         # 3^2 * 8
    assert api.Config.from_args([]).spp == 225                        # defaults: -t=1 -s=250 -> 15^2
    assert api.Config.from_args(["-t=10", "-s=1000"]).spp == 1000


def test_flag_parsing_matches_reference_regex(api):
    c = api.Config.from_args(["scenes/x", "--width=640", "-r=2", "--max-depth=7", "--light-bias=0.5", "-c=1,2,3",
                              "--unknown=1", "-bogus", "--gpu", "-w", "-a=2.8=1"]).raw
    assert c.scene_name == b"scenes/x"
    assert c.scene.has_output_width and c.scene.output_width == 640
    assert c.scene.has_aspect_ratio and c.scene.aspect_ratio == 2.0
    assert c.max_depth == 7 and c.light_bias == 0.5 and c.use_gpu == 1
    assert list(c.scene.camera_pos) == [1.0, 2.0, 3.0]
    assert not c.scene.has_f_number          # "-a=2.8=1" does not match ^-([^=\s]+)=([^=\s]+)$
    with pytest.raises(api.RtgpuError):
        api.Config.from_args(["--light-bias=1.5"])   # assert!(0 <= bias <= 1) in the reference
    with pytest.raises(api.RtgpuError):
        api.Config.from_args(["-w=abc"])


def test_camera_init_kat(api):
    # scenes/cornell: pos (277.5,277.5,-800) -> target z=0, focal 33, 600x600 (camera.rs:86-130)
    scene, cam, prm = api.load([api.scene_path("cornell")])
    assert (cam.width, cam.height) == (600, 600)
    h = 24.0 / 33.0
    vh = 800.0 * h
    assert cam.aperture_radius < 0
    # w = (0,0,-1); u = up x w = (-1,0,0); v = w x u = (0,1,0)
    assert np.allclose(list(cam.basis_u), [-1, 0, 0]) and np.allclose(list(cam.basis_v), [0, 1, 0])
    assert np.allclose(list(cam.pixel_delta_u), [-vh / 600, 0, 0])
    assert np.allclose(list(cam.pixel_delta_v), [0, -vh / 600, 0])
    fp = np.array([277.5, 277.5, 0.0]) - np.array([-vh, 0, 0]) / 2 - np.array([0, -vh, 0]) / 2 + np.array([-vh / 600, -vh / 600, 0]) * 0.5
    assert np.allclose(list(cam.first_pixel), fp)
    # height = trunc(width / aspect), aperture = (f/1000)/N
    scene, cam, prm = api.load([api.scene_path("light_test"), "-w=1001"])
    assert cam.height == int(1001 / 1.5)
    assert math.isclose(cam.aperture_radius, (70.0 / 1000.0) / 4.0)


def test_pitched_camera_keeps_unnormalised_basis(api):
    # u = v_up x w is not normalised (camera.rs:101-104)
    scene, cam, prm = api.load([api.scene_path("tonemap_test")])
    u = np.array(list(cam.basis_u))
    d = np.array([0, 30, 15.75]); w = d / np.linalg.norm(d)
    assert math.isclose(np.linalg.norm(u), np.linalg.norm(np.cross([0, 1, 0], w)), rel_tol=1e-12)
    assert np.linalg.norm(u) < 0.5


@pytest.mark.parametrize("name,objs,meshes,mats,imgs,noises", [
    ("cornell", 17, 0, 6, 0, 0), ("light_test", 7, 1, 4, 0, 0), ("perlin", 7, 1, 2, 0, 1),
    ("texture_test", 8, 1, 4, 3, 0), ("tonemap_test", 19, 0, 13, 0, 0), ("earth", 5, 0, 2, 1, 0), ("test", 5, 0, 2, 0, 0)])
def test_dsl_scenes_load(api, name, objs, meshes, mats, imgs, noises):
    scene, cam, prm = api.load([api.scene_path(name)])
    d = scene.desc.contents
    assert (d.n_objects, d.n_meshes, d.n_materials, d.n_images, d.n_noises) == (objs, meshes, mats, imgs, noises)
    assert scene.warnings == ""
    assert d.objects[d.world_root].kind == 0 and d.objects[d.lights_root].kind == 0


def test_dsl_transform_order_and_inverse(api, tmp_path):
    # transform.rs:44-92: ops apply left to right (transform = Op * transform; inv = inv * Op^-1)
    p = tmp_path / "s"
    p.write_text("m: lambertian (constant 1,1,1)\ns: sphere 0,0,0 1 $m\nt: transform $s t=1,2,3 ry=90 s=2\nworld: list $t\nlights: list\n")
    scene, cam, prm = api.load([str(p)])
    d = scene.desc.contents
    tr = [d.objects[i] for i in range(d.n_objects) if d.objects[i].kind == 3][0]
    M = np.array(list(tr.transform)).reshape(4, 4); Mi = np.array(list(tr.inv_transform)).reshape(4, 4)
    c, s = math.cos(math.pi / 2), math.sin(math.pi / 2)
    T = np.eye(4); T[:3, 3] = [1, 2, 3]
    R = np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]])
    S = np.diag([2, 2, 2, 1.0])
    assert np.allclose(M, S @ R @ T)
    assert np.allclose(M @ Mi, np.eye(4), atol=1e-12)


def test_dsl_errors_are_warnings_and_missing_world_fails(api, tmp_path):
    p = tmp_path / "s"
    p.write_text("m: lambertian (constant 1,1,1)\nbad line without colon\nx: frobnicate 1 2\ns: sphere 0,0,0 1 $nope\n"
                 "s2: sphere 0,0,0 1 $m\nworld: list $s2\nlights: list\n")
    scene, cam, prm = api.load([str(p)])
    assert scene.warnings.count("Warning") == 3
    q = tmp_path / "q"
    q.write_text("m: lambertian (constant 1,1,1)\n")
    with pytest.raises(api.RtgpuError, match="No world/lights"):
        api.load([str(q)])
    with pytest.raises(api.RtgpuError):
        api.load([str(tmp_path / "missing_file")])


def test_obj_loader(api, tmp_path):
    # obj.rs: v/vt/vn and v//vn, negative indices, vn normalised, unknown commands skipped
    (tmp_path / "m.obj").write_text("o x\nv 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nvt 0 0\nvt 1 0\nvt 0 1\nvn 0 0 2\ns 1\n"
                                    "f 1/1/1 2/2/1 3/3/1\nf -1//1 -3//-1 2//1\n")
    (tmp_path / "s").write_text("m: lambertian (constant 1,1,1)\no: mesh m.obj $m\nworld: list $o\nlights: list\n")
    scene, cam, prm = api.load([str(tmp_path / "s")])
    m = scene.desc.contents.meshes[0]
    assert (m.n_positions, m.n_normals, m.n_uvs, m.n_triangles) == (4, 1, 3, 2)
    assert [m.normals[i] for i in range(3)] == [0.0, 0.0, 1.0]
    assert [m.tri_pos[i] for i in range(6)] == [0, 1, 2, 3, 1, 1]
    assert [m.tri_uv[i] for i in range(6)] == [0, 1, 2, -1, -1, -1]


def test_obj_loader_chunked_parse_is_split_invariant(api, tmp_path, monkeypatch):
    """The OBJ text is parsed in parallel chunks (cut at line boundaries, merged in file order). Whatever the
    split, the mesh must equal the sequential parse — including negative indices, which are relative to the
    records seen so far in the WHOLE file (obj.rs:62-75)."""
    rng = np.random.default_rng(5)
    lines, nv, nt, nn = [], 0, 0, 0
    for blk in range(400):
        for _ in range(int(rng.integers(3, 9))):
            lines.append("v %.9g %.9g %.9g" % tuple(rng.normal(size=3))); nv += 1
        for _ in range(int(rng.integers(1, 4))):
            lines.append("vt %.6g %.6g" % tuple(rng.random(2))); nt += 1
        for _ in range(int(rng.integers(1, 3))):
            lines.append("vn %.6g %.6g %.6g" % tuple(rng.normal(size=3) + 0.1)); nn += 1
        lines.append("# block %d" % blk)
        for _ in range(int(rng.integers(2, 7))):
            if rng.random() < 0.5:   # relative indices
                c = ["%d/%d/%d" % (-int(rng.integers(1, nv + 1)), -int(rng.integers(1, nt + 1)), -int(rng.integers(1, nn + 1))) for _ in range(3)]
            elif rng.random() < 0.5:  # absolute indices
                c = ["%d/%d/%d" % (int(rng.integers(1, nv + 1)), int(rng.integers(1, nt + 1)), int(rng.integers(1, nn + 1))) for _ in range(3)]
            else:                     # no uvs
                c = ["%d//%d" % (-int(rng.integers(1, nv + 1)), int(rng.integers(1, nn + 1))) for _ in range(3)]
            lines.append("f " + " ".join(c))
    (tmp_path / "m.obj").write_text("\r\n".join(lines) + "\n")
    (tmp_path / "s").write_text("m: lambertian (constant 1,1,1)\no: mesh m.obj $m\nworld: list $o\nlights: list\n")

    def parse(threads):
        monkeypatch.setenv("RTGPU_OBJ_THREADS", str(threads))
        scene, _, _ = api.load([str(tmp_path / "s")])
        m = scene.desc.contents.meshes[0]
        arr = lambda p, n, t: np.ctypeslib.as_array(p, shape=(n,)).astype(t).copy()
        return (arr(m.positions, m.n_positions * 3, np.float64), arr(m.normals, m.n_normals * 3, np.float64),
                arr(m.uvs, m.n_uvs * 2, np.float64), arr(m.tri_pos, m.n_triangles * 3, np.int64),
                arr(m.tri_nrm, m.n_triangles * 3, np.int64), arr(m.tri_uv, m.n_triangles * 3, np.int64))

    ref = parse(1)
    assert len(ref[0]) == nv * 3 and len(ref[3]) > 3000
    for threads in (2, 7, 33):
        got = parse(threads)
        for a, b in zip(ref, got):
            assert np.array_equal(a, b)
    # an error is reported with its line number in the whole file
    (tmp_path / "m.obj").write_text("\n".join(lines[:900] + ["v 1 2"] + lines[900:]) + "\n")
    monkeypatch.setenv("RTGPU_OBJ_THREADS", "5")
    (tmp_path / "s").write_text("m: lambertian (constant 1,1,1)\no: mesh m.obj $m\nsp: sphere 0,0,0 1 $m\nworld: list $sp\nlights: list\n")
    scene, _, _ = api.load([str(tmp_path / "s")])   # the DSL turns the failed object into a warning (scene.rs:127-135)
    assert ":901:" in scene.warnings


def test_default_scene_is_seeded_and_has_bvh(api):
    a, _, _ = api.load([])
    b, _, _ = api.load([])
    c, _, _ = api.load(["--scene-seed=7"])
    da, db, dc = a.desc.contents, b.desc.contents, c.desc.contents
    assert da.n_objects == db.n_objects
    pa = [tuple(da.objects[i].p0) for i in range(da.n_objects) if da.objects[i].kind == 5]
    pb = [tuple(db.objects[i].p0) for i in range(db.n_objects) if db.objects[i].kind == 5]
    pc = [tuple(dc.objects[i].p0) for i in range(dc.n_objects) if dc.objects[i].kind == 5]
    assert pa == pb and pa != pc
    assert sum(1 for i in range(da.n_objects) if da.objects[i].kind == 1) > 100   # object BVH nodes (bvh.rs)
    assert any(da.objects[i].kind == 5 and da.objects[i].p1[0] < 0 for i in range(da.n_objects))  # hollow glass (Q15)


def test_png_roundtrip(api, tmp_path):
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (17, 23, 3), dtype=np.uint8)
    path = str(tmp_path / "o.png")
    api.save_png(path, img)
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(path)), img)
    (tmp_path / "s").write_text("t: image o.png\nm: lambertian $t\ns: sphere 0,0,0 1 $m\nworld: list $s\nlights: list\n")
    scene, _, _ = api.load([str(tmp_path / "s")])
    im = scene.desc.contents.images[0]
    got = np.ctypeslib.as_array(C.cast(im.data, C.POINTER(C.c_uint8)), shape=(17, 23, 3))
    assert np.array_equal(got, img)
