"""`model:<file>.glb` scenes: the AssimpLoader rules (src/loaders/assimp.rs:29-203) restated for binary
glTF. assets/models/test.glb is the reference's own models/test.glb (Blender export: plane, Suzanne, an emissive cube with
KHR_materials_emissive_strength, one perspective camera)."""
import json
import os
import struct

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GLB = os.path.join(ROOT, "assets", "models", "test.glb")


def _gltf_json():
    b = open(GLB, "rb").read()
    n = struct.unpack("<I", b[12:16])[0]
    return json.loads(b[20:20 + n])


def test_glb_scene_tables_follow_the_assimp_loader_rules(api):
    from rust_raytracer_b200 import _abi as abi
    scene, cam, prm = api.load(["model:" + GLB, "-w=120"])
    d = scene.desc.contents
    j = _gltf_json()
    # one TriangleMesh per primitive, one Transform(ObjectList) per node, an identity ROOT over the four scene roots
    kinds = [d.objects[i].kind for i in range(d.n_objects)]
    assert kinds.count(abi.OBJ_MESH) == 3 and kinds.count(abi.OBJ_TRANSFORM) == len(j["nodes"]) + 1
    assert [d.meshes[i].n_triangles for i in range(d.n_meshes)] == [2, 968, 12]
    root = d.objects[d.world_root]
    assert root.kind == abi.OBJ_TRANSFORM and list(root.transform) == [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]
    # materials (assimp.rs:143-189): two Glossy (base colour factor, roughness 0.5, ior 1.5), one Emissive = factor x strength
    mats = [d.materials[i] for i in range(d.n_materials)]
    assert [m.kind for m in mats] == [abi.MAT_GLOSSY, abi.MAT_GLOSSY, abi.MAT_EMISSIVE] and mats[0].ior == 1.5
    assert d.textures[mats[0].roughness].value[0] == 0.5
    assert np.allclose(list(d.textures[mats[0].albedo].value)[:3], j["materials"][0]["pbrMetallicRoughness"]["baseColorFactor"][:3])
    em = np.array(j["materials"][2]["emissiveFactor"]) * j["materials"][2]["extensions"]["KHR_materials_emissive_strength"]["emissiveStrength"]
    assert np.allclose(list(d.textures[mats[2].albedo].value)[:3], em, rtol=1e-6)
    # the emissive mesh's proxy light sphere (assimp.rs:123-129): centre = sum of the node chain's translations, radius =
    # the smallest vertex distance from the mesh origin (a unit-ish cube: its face centres ... its vertices here)
    lights = d.objects[d.lights_root]
    assert lights.kind == abi.OBJ_LIST and lights.n_children == 1
    sph = d.objects[d.children[lights.first_child]]
    assert sph.kind == abi.OBJ_SPHERE and sph.material == 2
    assert np.allclose(list(sph.p0)[:3], j["nodes"][3]["translation"])
    cube = d.meshes[2]
    pos = np.ctypeslib.as_array(cube.positions, shape=(cube.n_positions, 3))
    assert sph.p1[0] == np.sqrt((pos ** 2).sum(axis=1)).min()
    # the sphere is in `lights` only, not in the world
    world_kinds = set()
    def walk(i):
        o = d.objects[i]; world_kinds.add(o.kind)
        for k in range(o.n_children): walk(d.children[o.first_child + k])
    walk(d.world_root)
    assert abi.OBJ_SPHERE not in world_kinds
    # node transform = T * R * S in float precision; the Suzanne node's matrix times its inverse is the identity
    tr = [d.objects[i] for i in range(d.n_objects) if d.objects[i].kind == abi.OBJ_TRANSFORM]
    suz = [t for t in tr if np.allclose([t.transform[3], t.transform[7], t.transform[11]], j["nodes"][2]["translation"])][0]
    m, inv = np.array(list(suz.transform)).reshape(4, 4), np.array(list(suz.inv_transform)).reshape(4, 4)
    assert np.allclose(m @ inv, np.eye(4), atol=1e-12) and np.allclose(np.linalg.det(m[:3, :3]), 1.0, atol=1e-6)
    # v is flipped by the importer; indices are shared by positions, normals and uvs
    suzm = d.meshes[1]
    assert suzm.n_positions == suzm.n_normals == suzm.n_uvs
    tp = np.ctypeslib.as_array(suzm.tri_pos, shape=(suzm.n_triangles * 3,))
    tn = np.ctypeslib.as_array(suzm.tri_nrm, shape=(suzm.n_triangles * 3,))
    tu = np.ctypeslib.as_array(suzm.tri_uv, shape=(suzm.n_triangles * 3,))
    assert np.array_equal(tp, tn) and np.array_equal(tp, tu)
    # camera (assimp.rs:41-51): node-local position (0,0,0) -> (0,0,-1), aspect 1.5, focal = 18 / tan(hfov / 2) = 24 mm
    sc = scene.cfg.raw.scene
    assert list(sc.camera_pos) == [0, 0, 0] and list(sc.camera_target) == [0, 0, -1]
    assert sc.aspect_ratio == 1.5 and abs(sc.focal_length - 24.0) < 1e-5
    assert (cam.width, cam.height) == (120, 80)
    # the command line overrides the model's camera (SceneConfig::merge, assimp.rs:53)
    scene2, cam2, _ = api.load(["model:" + GLB, "-c=7,5,7", "-f=50"])
    assert list(cam2.position) == [7, 5, 7] and scene2.cfg.raw.scene.focal_length == 50


def test_glb_loader_refuses_what_the_reference_cannot_load(api, tmp_path):
    with pytest.raises(api.RtgpuError) as e:
        api.load(["model:" + str(tmp_path / "missing.glb")])
    assert e.value.code == -5
    (tmp_path / "a.fbx").write_bytes(b"Kaydara FBX Binary  \x00" + b"\x00" * 64)
    with pytest.raises(api.RtgpuError) as e:
        api.load(["model:" + str(tmp_path / "a.fbx")])
    assert e.value.code == -4 and "binary glTF" in str(e.value)
    # a textured material: embedded image, which the reference's `<dir>/<name>.<ext>` lookup cannot open either
    raw = open(GLB, "rb").read()
    n = struct.unpack("<I", raw[12:16])[0]
    j = json.loads(raw[20:20 + n])
    j["materials"][0]["pbrMetallicRoughness"]["baseColorTexture"] = {"index": 0}
    js = json.dumps(j).encode()
    js += b" " * (-len(js) % 4)
    rest = raw[20 + n:]
    out = b"glTF" + struct.pack("<II", 2, 12 + 8 + len(js) + len(rest)) + struct.pack("<I", len(js)) + b"JSON" + js + rest
    (tmp_path / "tex.glb").write_bytes(out)
    with pytest.raises(api.RtgpuError) as e:
        api.load(["model:" + str(tmp_path / "tex.glb")])
    assert "embedded" in str(e.value)


def test_glb_scene_renders_with_the_oracle_and_the_host_simulated_traversal(api, oracle):
    """the loaded tables are a valid scene for both consumers: the CPU oracle renders it (light reaches the plane through
    the proxy sphere's light sampling), and first hits of the flattened wide BVH agree with the oracle's octree walk."""
    import ctypes as C
    import subprocess
    from util import primary_rays
    scene, cam, prm = api.load(["model:" + GLB, "-w=60", "-s=16", "-t=4", "-c=7.36,4.96,6.93", "-l=2.5,1,4.5"])
    o = oracle.OracleScene(scene)
    img, _ = o.render(cam, prm, n_threads=4)
    assert np.isfinite(img).all() and img[..., 0].mean() > 0.5 and img[..., 0].mean() > img[..., 2].mean()   # orange light
    out = os.path.join(ROOT, "build", "libhostsim.so")
    if not os.path.exists(out):
        pytest.skip("hostsim library not built yet (tests/test_hostsim.py builds it)")
    sim = C.CDLL(out)
    h = C.c_void_p()
    assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0
    rays, _ = primary_rays(oracle, cam, 20_000, seed=4)
    ho = o.trace_primary(rays)
    n = len(rays)
    inst = np.zeros(n, np.int32); prim = np.zeros(n, np.int32); t = np.zeros(n)
    sim.hostsim_closest(h, rays.ctypes.data_as(C.POINTER(C.c_double)), C.c_int64(n), 1, inst.ctypes.data_as(C.POINTER(C.c_int32)),
                        prim.ctypes.data_as(C.POINTER(C.c_int32)), t.ctypes.data_as(C.POINTER(C.c_double)))
    hit = ho["object"] >= 0
    assert hit.mean() > 0.5
    assert np.array_equal(np.where(hit, ho["instance"], -1), inst) and np.array_equal(np.where(hit, ho["primitive"], 0), prim)
    assert np.array_equal(ho["t"][hit], t[hit])
    sim.hostsim_scene_destroy(h)
