"""Device logic on the CPU: the traversal / intersection headers (csrc/device/*.cuh) are compiled
for the host by tests/hostsim (test infrastructure only, never linked into librtgpu.so) so that
the 8-wide BVH builder and both trace modes can be checked against the oracle without a GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from util import primary_rays

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def sim():
    out = os.path.join(ROOT, "build", "libhostsim.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    dev = os.path.join(ROOT, "rust_raytracer_b200", "csrc", "device")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include"),
                           "-I" + dev, "-x", "c++", os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp"),
                           os.path.join(dev, "flatten.cpp"), os.path.join(dev, "bvh.cpp"), "-o", out])
    lib = C.CDLL(out)
    lib.hostsim_last_error.restype = C.c_char_p
    return lib


def _closest(sim, h, rays, exact):
    n = rays.shape[0]
    inst = np.zeros(n, np.int32); prim = np.zeros(n, np.int32); t = np.zeros(n)
    sim.hostsim_closest(h, rays.ctypes.data_as(C.POINTER(C.c_double)), C.c_int64(n), exact, inst.ctypes.data_as(C.POINTER(C.c_int32)),
                        prim.ctypes.data_as(C.POINTER(C.c_int32)), t.ctypes.data_as(C.POINTER(C.c_double)))
    return inst, prim, t


@pytest.mark.parametrize("args", [["SCENE:cornell"], ["SCENE:light_test"], ["SCENE:perlin"], ["-w=600"], ["cornell_smoke"]],
                         ids=lambda a: a[0])
def test_wide_bvh_traversal_matches_oracle(api, oracle, sim, args):
    args = [api.scene_path(a[6:]) if a.startswith("SCENE:") else a for a in args]
    scene, cam, prm = api.load(args)
    o = oracle.OracleScene(scene)
    h = C.c_void_p()
    assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0, sim.hostsim_last_error()
    info = (C.c_int64 * 4)()
    sim.hostsim_bvh_info(h, info)
    assert info[2] <= 32                                     # BVH depth fits the traversal stack
    rays, _ = primary_rays(oracle, cam, 60_000, seed=2)
    ho = o.trace_primary(rays)
    d = scene.desc.contents
    finite = (ho["object"] >= 0) & (ho["t"] < 1e300)
    vol = np.array([d.objects[i].kind == 7 if i >= 0 else False for i in ho["object"]])
    finite &= ~vol                                           # stochastic volumes are not part of the closest-hit search
    oi = np.where(finite, ho["instance"], -1); op = np.where(finite, ho["primitive"], 0)
    inst, prim, t = _closest(sim, h, rays, 1)
    if not vol.any():
        assert np.array_equal(oi, inst) and np.array_equal(op, prim)
        assert np.array_equal(ho["t"][finite], t[finite])    # bit-equal f64 t in exact mode
    inst, prim, t = _closest(sim, h, rays, 0)
    agree = (oi == inst) & (op == prim)
    if not vol.any():
        assert agree.mean() > 0.9995                         # fp32 mode: only silhouette-edge rays may differ
        both = finite & agree
        assert np.allclose(ho["t"][both], t[both], rtol=2e-5)
    sim.hostsim_scene_destroy(h)


def test_parallel_build_is_thread_invariant(api, sim, monkeypatch):
    """The binary SAH build hands big subtrees to threads and the wide collapse runs per-subtree jobs; the
    resulting node array, leaf order and leaf records must not depend on thread count or split sizes."""
    sim.hostsim_bvh_hash.restype = C.c_uint64
    scene, cam, prm = api.load(["-w=600"])   # default scene: Suzanne (15.7 k triangles) + ~440 spheres
    hashes = []
    for threads, par_min in ((1, 1 << 30), (2, 64), (8, 256), (5, 4096)):
        monkeypatch.setenv("RTGPU_BUILD_THREADS", str(threads))
        monkeypatch.setenv("RTGPU_BUILD_PAR_MIN", str(par_min))
        h = C.c_void_p()
        assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0, sim.hostsim_last_error()
        hashes.append(sim.hostsim_bvh_hash(h))
        sim.hostsim_scene_destroy(h)
    assert len(set(hashes)) == 1, hashes


def test_quantised_boxes_are_conservative(api, sim):
    """Every child box of every wide node must contain the primitives below it (checked through
    traversal: brute force over all primitives finds no closer hit than the BVH does)."""
    scene, cam, prm = api.load([api.scene_path("light_test")])
    h = C.c_void_p()
    assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0
    rng = np.random.default_rng(0)
    n = 20_000
    o = rng.normal(size=(n, 3)) * 4
    tgt = rng.uniform(-1.2, 1.2, size=(n, 3)) + [0, 0, -1.5]
    rays = np.concatenate([o, tgt - o], 1)
    inst, prim, t = _closest(sim, h, rays, 1)
    import oracle as orc
    ho = orc.OracleScene(scene).trace_primary(rays)
    fin = ho["object"] >= 0
    assert np.array_equal(np.where(fin, ho["primitive"], 0), prim)
    assert np.array_equal(ho["t"][fin], t[fin])
    sim.hostsim_scene_destroy(h)


@pytest.mark.parametrize("name", ["COVERAGE", "cornell", "light_test"])
def test_bounce_hook_matches_oracle_on_the_host(api, oracle, sim, name):
    """bounce_hook_one (hooks.cuh) is the function k_bounce_hook runs per ray; compiled for the host it is compared
    here with the oracle's bounce for the same rays and the same injected draws — in particular the position of
    Volume::test's free-flight draw in the stream (volume.rs:57: inside world.test, per volume in list order, only when
    the boundary segment seen at that moment is not empty) and every material / texture branch of coverage_test.
    The GPU run of the same comparison is tests/test_gpu_parity.py::test_bounce_parity_injected."""
    from util import load_coverage_scene
    from rust_raytracer_b200 import _abi as abi
    scene, cam, prm = load_coverage_scene(api) if name == "COVERAGE" else api.load([api.scene_path(name)])
    o = oracle.OracleScene(scene)
    h = C.c_void_p()
    assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0, sim.hostsim_last_error()
    n, n_xi = 40_000, 14
    rays, _ = primary_rays(oracle, cam, n, seed=5)
    xi = np.random.default_rng(7).uniform(0.02, 0.98, (n, n_xi))
    bo = o.bounce(rays, xi, 0.25)
    out = (abi.Bounce * n)()
    sim.hostsim_bounce(h, C.c_double(0.25), rays.ctypes.data_as(C.POINTER(C.c_double)), xi.ctypes.data_as(C.POINTER(C.c_double)),
                       n_xi, C.c_int64(n), out)
    bg = np.frombuffer(out, dtype=api.BOUNCE_DTYPE, count=n).copy()
    sim.hostsim_scene_destroy(h)
    same = bo["kind"] == bg["kind"]
    assert same.mean() > 0.9995, same.mean()
    if name == "COVERAGE":
        medium = (bo["kind"] == 1) & np.isclose(bo["scattering_pdf"], 1.0 / (4.0 * np.pi))
        assert medium.mean() > 0.10                       # more than a tenth of the compared bounces are volume hits
        assert (bg["kind"][medium] == 1).all()
        assert np.allclose(bg["next_origin"][medium], bo["next_origin"][medium], rtol=1e-9, atol=1e-9)   # same free-flight distance
        assert set(np.unique(bo["kind"])) >= {1, 2, 4}
    ok = same & (np.isfinite(bo["weight"]).all(axis=1) & (bo["pdf"] > 1e-6) | (bo["kind"] != 1))
    off = ~np.isclose(bo["emission"][ok], bg["emission"][ok], rtol=1e-4, atol=1e-5).all(axis=1)
    assert off.mean() < 1e-4
    cont = ok & np.isin(bo["kind"], [1, 2])
    d_o = bo["next_dir"][cont] / np.linalg.norm(bo["next_dir"][cont], axis=1, keepdims=True)
    d_g = bg["next_dir"][cont] / np.linalg.norm(bg["next_dir"][cont], axis=1, keepdims=True)
    assert np.quantile(np.sum(d_o * d_g, axis=1), 0.001) > 1 - 1e-4
    rel = np.abs(bo["weight"][cont] - bg["weight"][cont]) / (np.abs(bo["weight"][cont]) + 1e-3)
    assert np.quantile(rel, 0.999) < 5e-3


def test_batch_plan_rules(sim):
    """batch_plan.h (api.cu: enqueue_render): equal batches of whole samples, a whole number per lane, small jobs split over the lanes,
    jobs of between one and two batches per lane as one larger batch per lane — and the memory budget respected throughout."""
    def plan(total, npix, batch_mi=40, lanes=4, bpp=280.0, budget=107e9):
        out = (C.c_uint64 * 3)()
        rc = sim.hostsim_plan_batches(C.c_uint64(total), C.c_uint64(npix), C.c_uint64(batch_mi << 20), lanes, C.c_double(bpp), C.c_double(budget), out)
        return rc, int(out[0]), int(out[1]), int(out[2])
    px = 1200 * 1200
    assert plan(1000, px) == (0, 28, 36, 4)                 # the headline job: 36 equal batches of 28 samples (at most 29 fit 40 Mi paths)
    assert plan(125, px) == (0, 32, 4, 4)                   # one GPU's share at N = 8: 5 batches of 29 -> ONE batch of 32 per lane
    assert plan(250, px) == (0, 21, 12, 4)                  # 9 batches of 29 -> 12 equal ones, three per lane
    assert plan(90, 600 * 600) == (0, 23, 4, 4)             # BASELINE config 1 fits one batch: split into one per lane (>= 4 Mi paths each)
    assert plan(64, 96 * 96) == (0, 64, 1, 1)               # too small to split
    assert plan(1, px) == (0, 1, 1, 1)
    assert plan(1000, px, lanes=1) == (0, 29, 35, 1)        # one lane (profiling): nothing to balance
    # a tight budget halves the batches, then drops lanes; an impossible one fails
    rc, spb, nb, nl = plan(1000, px, budget=4 * 8e9)
    assert rc == 0 and nl == 4 and spb * px * 280.0 * nl <= 4 * 8e9 and spb < 28
    rc, spb, nb, nl = plan(1000, px, budget=1.5 * px * 280.0)
    assert rc == 0 and (spb, nl) == (1, 1)
    assert plan(1000, px, budget=0.5 * px * 280.0)[0] == -1
    # properties over a sweep: every sample in exactly one batch, no lane idle while another holds two more batches, budget kept
    rng = np.random.default_rng(5)
    for _ in range(3000):
        total = int(rng.integers(1, 5000)); npix = int(rng.integers(1, 4_000_000)); lanes = int(rng.integers(1, 7))
        batch_mi = int(rng.integers(1, 64)); budget = float(rng.uniform(2e9, 120e9))
        rc, spb, nb, nl = plan(total, npix, batch_mi, lanes, 280.0, budget)
        if rc != 0:
            assert npix * 280.0 > budget
            continue
        assert 1 <= nl <= lanes and nb >= nl and spb >= 1
        assert (nb - 1) * spb < total <= nb * spb
        assert nl * spb * npix * 280.0 <= budget
        assert spb * npix <= max(batch_mi << 20, npix) or nb == nl   # larger than the nominal batch only as one batch per lane


def test_broadcast_list_carries_constant_emission(api, sim):
    """flatten.cpp marks a broadcast-list primitive whose material is Emissive over a constant map (BCAST_EMIT_CONST, bit 8) and stores
    the emission in the .w lanes of its record: inline_emission (kernels.cuh) then needs no material / texture look-up. cornell: the
    ceiling light (emissive 15,15,15) is the only such entry; the glass ball is on the list as a sphere, not inward, not emissive."""
    scene, cam, prm = api.load([api.scene_path("cornell")])
    h = C.c_void_p()
    assert sim.hostsim_scene_create(scene.desc, C.byref(h)) == 0, sim.hostsim_last_error()
    flags, em = C.c_uint32(), (C.c_float * 3)()
    n = sim.hostsim_bcast_info(h, 0, C.byref(flags), em)
    assert 6 <= n <= 16
    emitters, spheres = [], 0
    for i in range(n):
        assert sim.hostsim_bcast_info(h, i, C.byref(flags), em) == n
        f = flags.value
        cls = (f >> 4) & 7
        if f & (1 << 8):                                   # BCAST_EMIT_CONST
            assert cls == 5 and not (f & (1 << 9))         # CLASS_EMISSIVE, not an inward sphere
            emitters.append(tuple(em))
        else:
            assert tuple(em) == (0.0, 0.0, 0.0)
        spheres += bool(f & (1 << 3))                      # PRIM_SPHERE
    assert emitters == [(15.0, 15.0, 15.0)] and spheres == 1
    sim.hostsim_scene_destroy(h)
