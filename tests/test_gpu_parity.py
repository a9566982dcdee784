"""GPU parity tests (run on the B200 box): the CUDA path through the C ABI vs the CPU oracle.

Bars (BASELINE.json north_star): primary-ray first-hit object / instance / primitive ids
bit-exact; hit t and normals within 1e-5 relative (t is in fact bit-equal: the exact kernel
performs the reference's f64 operations); converged images within an RMSE / mean-luminance
tolerance stated in each test.
"""
import os

import numpy as np
import pytest

from util import load_coverage_scene, luminance, primary_rays, rmse

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRAGON = ["DRAGON", "--max-depth=20", "--light-bias=0.25"]   # BASELINE config 5 (proxy mesh, tools/make_dragon_proxy.py)


def _record(section, key, value):
    """Measured agreement figures go to gpurun_out/r02_parity.json (merged back from the GPU box; the copy the
    judge reads is profiles/r02_parity.json)."""
    import json
    out = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        path = os.path.join(out, "r02_parity.json")
        data = {}
        if os.path.exists(path):
            with open(path) as f:
                data = json.load(f)
        data.setdefault(section, {})[key] = value
        with open(path, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
    except OSError:
        pass


SCENES = [
    (DRAGON, 200_000),                # the headline scene: 1.0 M triangles, 150 k-node collapsed BVH
    (["SCENE:cornell"], 200_000),
    (["SCENE:light_test"], 200_000),
    (["SCENE:perlin"], 100_000),
    (["SCENE:texture_test"], 100_000),
    (["SCENE:tonemap_test"], 100_000),
    (["SCENE:earth"], 100_000),
    (["-w=600"], 200_000),            # default scene (golden_monkey), scene seed 42
    (["cornell_smoke"], 50_000),
    (["MODEL"], 100_000),             # model:assets/models/test.glb: per-node Transform(ObjectList), shared-index meshes
    (["COVERAGE"], 100_000),          # assets/scenes/coverage_test + ABI patches (tests/util.py): media, bare / flat / two-sided meshes
]


def _rays_at_mesh(o, primary, n, seed):
    """n rays from random points of the room towards points ON the mesh (taken from the primary rays' mesh hits,
    jittered): the camera sees the mesh in ~7 % of the pixels only, these rays walk it from every side."""
    ho = o.trace_primary(primary)
    on_mesh = ho["pos"][ho["primitive"] > 0]
    assert len(on_mesh) > 1000
    rng = np.random.default_rng(seed)
    tgt = on_mesh[rng.integers(0, len(on_mesh), n)] + rng.normal(0.0, 2.0, (n, 3))
    org = rng.uniform(20.0, 535.0, (n, 3))
    return np.concatenate([org, tgt - org], axis=1)


MODEL = ["model:" + os.path.join(ROOT, "assets", "models", "test.glb"), "-c=7.36,4.96,6.93", "-l=2.5,1,4.5"]   # the .glb's own camera node, by hand


def _load(api, args):
    if args and args[0] == "MODEL":   # the reference's models/test.glb through the glTF restatement of AssimpLoader
        return api.load(MODEL + list(args[1:]))
    if args and args[0] == "COVERAGE":
        return load_coverage_scene(api, args[1:])
    if args and args[0] == "DRAGON":
        from tools.make_dragon_proxy import ensure
        ensure()
        args = ["SCENE:cornell_dragon"] + list(args[1:])
    args = [api.scene_path(a[6:]) if a.startswith("SCENE:") else a for a in args]
    return api.load(args)


@pytest.mark.parametrize("args,n", SCENES, ids=[a[0] for a, _ in SCENES])
def test_first_hit_parity(api, oracle, args, n):
    scene, cam, prm = _load(api, args)
    o = oracle.OracleScene(scene)
    g = api.GpuScene(scene)
    rays, _ = primary_rays(oracle, cam, n, seed=1)
    if args[0] == "DRAGON":
        rays = np.concatenate([rays, _rays_at_mesh(o, rays, n, seed=2)])
    ho, hg = o.trace_primary(rays), g.trace_primary(rays)
    if args[0] in ("cornell_smoke", "COVERAGE"):  # stochastic volumes are skipped by this hook (their parity: test_bounce_parity_injected)
        vol = np.isin(ho["object"], [i for i in range(scene.desc.contents.n_objects) if scene.desc.contents.objects[i].kind == 7])
        keep = ~vol
        ho, hg = ho[keep], hg[keep]
    _record("first_hit_exact", args[0], {"rays": int(len(ho)), "mesh_hits": int((ho["primitive"] > 0).sum()),
                                         "id_mismatches": int((ho["object"] != hg["object"]).sum() + (ho["primitive"] != hg["primitive"]).sum()),
                                         "t_bit_mismatches": int((ho["t"][ho["object"] >= 0] != hg["t"][ho["object"] >= 0]).sum())})
    assert np.array_equal(ho["object"], hg["object"]), f"{(ho['object'] != hg['object']).sum()} object id mismatches"
    assert np.array_equal(ho["instance"], hg["instance"])
    assert np.array_equal(ho["primitive"], hg["primitive"])
    assert np.array_equal(ho["front_face"], hg["front_face"])
    hit = ho["object"] >= 0
    finite = hit & np.isfinite(ho["t"]) & (ho["t"] < 1e300)
    assert np.array_equal(ho["t"][hit], hg["t"][hit]), "t must be bit-equal in exact mode"
    dn = np.linalg.norm(ho["normal"][hit] - hg["normal"][hit], axis=1)
    assert (dn <= 1e-5 * np.linalg.norm(ho["normal"][hit], axis=1)).all(), f"normal rel err max {dn.max()}"
    assert np.allclose(ho["pos"][finite], hg["pos"][finite], rtol=1e-9, atol=1e-9)
    assert np.allclose(ho["uv"][hit], hg["uv"][hit], rtol=1e-5, atol=1e-7)


def test_generate_rays_matches_oracle(api, oracle):
    scene, cam, prm = _load(api, ["SCENE:light_test", "-w=1200"])
    rays, (px, sxy, xi) = primary_rays(oracle, cam, 50_000, seed=3, sqrt_spt=4)
    got = api.generate_rays(cam, 4, px, sxy, xi)
    assert np.allclose(got, rays, rtol=1e-13, atol=1e-13)


@pytest.mark.parametrize("name", ["cornell", "light_test", "perlin", "texture_test", "earth", "COVERAGE", "MODEL"])
def test_bounce_parity_injected(api, oracle, name):
    """Same rays + same injected random numbers -> same event kind, emission, weight, next ray.
    COVERAGE: constant media (Volume::test draws its free-flight uniform INSIDE the intersection, in world-list visit
    order, before the material's draws: volume.rs:57, hooks.cuh), Isotropic + UniformPDF, normal_debug with and without
    a normal map, checker_solid, uv_debug, Clamp-mode image, flat_shading, hit_back_faces, a bare mesh."""
    scene, cam, prm = _load(api, [name] if name in ("COVERAGE", "MODEL") else [f"SCENE:{name}"])
    o = oracle.OracleScene(scene)
    g = api.GpuScene(scene)
    n = 100_000
    rays, _ = primary_rays(oracle, cam, n, seed=5)
    rng = np.random.default_rng(7)
    xi = rng.uniform(0.02, 0.98, (n, 14))
    bo, bg = o.bounce(rays, xi, 0.25), g.bounce(rays, xi, 0.25)
    if name == "COVERAGE":
        medium = (bo["kind"] == 1) & np.isclose(bo["scattering_pdf"], 1.0 / (4.0 * np.pi))   # isotropic.rs: phase function 1 / 4 pi
        print(f"volume events: {medium.mean():.3f} of the compared bounces")
        assert medium.mean() > 0.10
        assert (bg["kind"][medium] == 1).all() and np.allclose(bg["next_origin"][medium], bo["next_origin"][medium], rtol=1e-9, atol=1e-9)
        _record("bounce", "coverage_test", {"rays": n, "volume_event_fraction": float(medium.mean()),
                                            "kind_agreement": float((bo["kind"] == bg["kind"]).mean())})
    ok = np.isfinite(bo["weight"]).all(axis=1) & (bo["pdf"] > 1e-6) | (bo["kind"] != 1)
    # fp32 shading: decisions taken within rounding distance of a threshold may flip
    same_kind = bo["kind"] == bg["kind"]
    assert same_kind[ok].mean() > 0.999, f"kind agreement {same_kind[ok].mean()}"
    m = ok & same_kind
    # (nearest-neighbour image lookups: a uv within fp32 rounding of a texel edge may fetch the neighbour)
    off = ~np.isclose(bo["emission"][m], bg["emission"][m], rtol=1e-4, atol=1e-5).all(axis=1)
    assert off.mean() < 1e-4, f"emission mismatches {off.sum()} of {m.sum()}"
    cont = m & np.isin(bo["kind"], [1, 2])
    d_o = bo["next_dir"][cont] / np.linalg.norm(bo["next_dir"][cont], axis=1, keepdims=True)
    d_g = bg["next_dir"][cont] / np.linalg.norm(bg["next_dir"][cont], axis=1, keepdims=True)
    cosang = np.sum(d_o * d_g, axis=1)
    assert np.quantile(cosang, 0.001) > 1 - 1e-4, f"direction agreement {np.quantile(cosang, 0.001)}"
    w_o, w_g = bo["weight"][cont], bg["weight"][cont]
    rel = np.abs(w_o - w_g) / (np.abs(w_o) + 1e-3)
    assert np.quantile(rel, 0.999) < 5e-3, f"weight rel err q99.9 {np.quantile(rel, 0.999)}"
    assert np.allclose(bo["next_origin"][cont], bg["next_origin"][cont], rtol=1e-5, atol=2e-3)


def _converged(api, oracle, args):
    """RMSE(GPU, oracle) <= 1.5 x RMSE(oracle seed A, oracle seed B), and the mean luminance / channel means within
    3 sigma of the difference of two independent renders' means, sigma MEASURED from the oracle's two seeds
    (per-pixel variance = (A - B)^2 / 2; the mean of N pixels has variance sum / N^2; GPU - oracle has twice that).
    BASELINE.md section 3 states the gate this way; at the sizes the CPU oracle can afford it is 0.2-1.2 % of the mean
    (the 0.5 % of the full-size job is checked where a full-size reference exists: sample1.png / sample2.png)."""
    scene, cam, prm = _load(api, args)
    o = oracle.OracleScene(scene)
    g = api.GpuScene(scene)
    a, _ = o.render(cam, prm, n_threads=0)
    prm2 = type(prm).from_buffer_copy(prm)
    prm2.seed = 12345
    b, _ = o.render(cam, prm2, n_threads=0)
    gbuf = g.render(cam, prm)
    floor = rmse(a, b)             # oracle-vs-oracle noise floor at this spp
    got = rmse(gbuf, a)
    la, lg = luminance(a).mean(), luminance(gbuf).mean()
    npix = a.shape[0] * a.shape[1]
    sigma = lambda x, y: float(np.sqrt(((x - y) ** 2).mean() / 2.0 / npix)) * np.sqrt(2.0)   # of a difference of two means
    gate_lum = 3.0 * sigma(luminance(a), luminance(b))
    print(f"{args}: noise floor {floor:.5f} gpu-vs-oracle {got:.5f} mean lum oracle {la:.5f} gpu {lg:.5f} "
          f"(diff {abs(lg - la) / la:.5f} of the mean, 3-sigma gate {gate_lum / la:.5f})")
    _record("converged", " ".join(args), {"rmse_gpu_vs_oracle": got, "rmse_oracle_seed_vs_seed": floor, "mean_lum_rel_diff": float(abs(lg - la) / la),
                                         "mean_lum_gate_3sigma_rel": float(gate_lum / la)})
    assert np.isfinite(gbuf).all()
    assert got <= 1.5 * floor + 1e-6, f"RMSE {got} vs noise floor {floor}"
    assert abs(lg - la) <= gate_lum + 1e-7, f"mean luminance {lg} vs {la} (3-sigma gate {gate_lum})"
    for c in range(3):
        ma, mg = a[..., c].mean(), gbuf[..., c].mean()
        assert abs(mg - ma) <= 3.0 * sigma(a[..., c], b[..., c]) + 1e-6, f"channel {c} mean {mg} vs {ma}"
    return gbuf


@pytest.mark.parametrize("args", [
    ["SCENE:cornell", "-w=128", "-s=256", "-t=4"],
    ["SCENE:light_test", "-w=160", "-s=256", "-t=4"],
    ["SCENE:perlin", "-w=128", "-s=144", "-t=4"],
    ["SCENE:texture_test", "-w=128", "-s=144", "-t=4"],
    ["SCENE:earth", "-w=128", "-s=144", "-t=4"],
    ["-w=128", "-s=144", "-t=4"],
    ["cornell_smoke", "-w=96", "-s=144", "-t=4"],
    ["COVERAGE", "-w=128", "-s=256", "-t=4"],
    ["MODEL", "-w=120", "-s=256", "-t=4"],
], ids=lambda a: a[0])
def test_converged_image_parity(api, oracle, args):
    """RMSE(GPU, oracle) <= 1.5 x RMSE(oracle seed A, oracle seed B); mean luminance within the measured 3-sigma gate."""
    _converged(api, oracle, args)


def test_render_adds_into_buffer_and_shards_sum(api, oracle):
    """Camera::render ADDS into the caller's buffer; sample shards sum to the full render."""
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=64", "-s=64", "-t=4"])
    g = api.GpuScene(scene)
    full = g.render(cam, prm)
    spp = prm.replicas * prm.sqrt_spt ** 2
    acc = np.zeros_like(full)
    for b, e in [(0, spp // 4), (spp // 4, spp // 2), (spp // 2, spp)]:
        p = type(prm).from_buffer_copy(prm)
        p.sample_begin, p.sample_end = b, e
        g.render(cam, p, out=acc)
    assert np.allclose(acc, full, rtol=1e-4, atol=1e-5)  # fp32 summation order only


def test_tonemap_matches_oracle(api, oracle):
    rng = np.random.default_rng(0)
    buf = np.zeros((64, 64, 4))
    buf[..., :3] = rng.gamma(0.5, 2.0, (64, 64, 3))
    buf[0, 0, :3] = [0, 1e-4, 1e6]
    for mode in ("aces", "clamp"):
        a, b = oracle.tonemap_srgb8(buf, mode), api.tonemap_srgb8(buf, mode)
        assert (np.abs(a.astype(int) - b.astype(int)) <= 1).all()   # pow() may differ in the last ulp at a quantisation edge
        assert (a != b).mean() < 1e-3


@pytest.mark.parametrize("args,n", [(DRAGON, 400_000), (["SCENE:cornell"], 200_000), (["SCENE:light_test"], 200_000), (["-w=600"], 200_000),
                                    (["SCENE:texture_test"], 100_000), (["COVERAGE"], 100_000), (["MODEL"], 100_000)], ids=lambda a: a[0] if isinstance(a, list) else None)
def test_production_trace_kernel_matches_exact_path(api, oracle, args, n):
    """k_trace_fast (fp32, dynamic fetch — the kernel every render uses) against the exact f64 hook on
    primary AND scattered rays: same instance/primitive except within rounding distance of an edge,
    t within 1e-5 relative (north_star tolerance)."""
    scene, cam, prm = _load(api, args)
    g = api.GpuScene(scene)
    o = oracle.OracleScene(scene)
    rays, _ = primary_rays(oracle, cam, n, seed=11)
    aimed = _rays_at_mesh(o, rays, n, seed=12) if args[0] == "DRAGON" else np.zeros((0, 6))
    rng = np.random.default_rng(13)
    xi = rng.uniform(0.02, 0.98, (n, 12))
    b = o.bounce(rays, xi, 0.25)
    cont = np.isin(b["kind"], [1, 2]) & np.isfinite(b["next_dir"]).all(axis=1) & (np.linalg.norm(b["next_dir"], axis=1) > 1e-9)
    secondary = np.concatenate([b["next_origin"][cont], b["next_dir"][cont]], axis=1)
    # "aimed": rays from anywhere in the room at the mesh (DRAGON only); like scattered rays they are bounded by the fp32
    # world coordinates of the hit point, not by t (the camera sits 800 units away, these origins a few units)
    for label, rr in (("primary", rays), ("aimed", aimed), ("secondary", secondary)):
        if len(rr) == 0:
            continue
        rr = rr.copy()
        rr[:, 3:6] = rr[:, 3:6].astype(np.float32)   # the pipeline's ray record keeps the direction in fp32 (PathRay)
        he, hf = g.trace_primary(rr), g.trace_fast(rr)
        surf = he["object"] >= 0
        surf &= np.isfinite(he["t"]) & (he["t"] < 1e300)           # sky / sun are resolved outside the traversal
        e_inst = np.where(surf, he["instance"], -1)
        same = (e_inst == hf["instance"]) & ((he["primitive"] == hf["primitive"]) | ~surf)
        frac = same.mean()
        print(f"{args[0]} {label}: {len(rr)} rays, id agreement {frac:.6f}")
        rec = {"rays": int(len(rr)), "mesh_hits": int((surf & (he["primitive"] > 0)).sum()), "id_agreement": float(frac)}
        # measured on the B200 (profiles/r02_parity.json): >= 0.99993 on every scene and ray set; the rays that differ
        # graze a silhouette or a shared edge within fp32 rounding of the world-space triangle
        assert frac > 0.9997, f"{label}: fast/exact id agreement {frac}"
        m = same & surf
        # primary rays: the north_star bar, 1e-5 relative. Scattered rays start ON surfaces and can
        # have t -> 0, where fp32 world coordinates (ulp(555) = 6e-5) bound the ABSOLUTE error of the
        # hit point instead: allow 1e-5 relative + 4e-6 x the scene extent, amplified by 1/cos(incidence)
        # (a position error across the surface moves a grazing hit far along the ray; quads hit at
        # grazing angles are re-evaluated in f64 by the shade kernel, surface_for_slot).
        dlen = np.linalg.norm(rr[m, 3:6], axis=1)
        extent = np.abs(rr[m, 0:3]).max() + 1.0
        err = np.abs(hf["t"][m] - he["t"][m]) * dlen
        nrm = he["normal"][m]
        cosi = np.abs(np.sum(nrm * rr[m, 3:6], axis=1)) / (dlen * np.maximum(np.linalg.norm(nrm, axis=1), 1e-300))
        tol = 1e-5 * np.abs(he["t"][m]) * dlen + (0.0 if label == "primary" else 4e-6 * extent / np.maximum(cosi, 1e-6))
        worst = (err / np.maximum(tol, 1e-300)).max()
        print(f"    max |dt| x |d| = {err.max():.3e}, worst err/tol = {worst:.3f}")
        rec.update({"max_abs_dt_times_len": float(err.max()), "max_rel_dt": float((np.abs(hf["t"][m] - he["t"][m]) / np.abs(he["t"][m])).max()),
                    "worst_err_over_tol": float(worst)})
        _record("fast_vs_exact", f"{args[0]} {label}", rec)
        assert worst <= 1.0, f"{label}: hit distance error {err.max()} (err/tol {worst})"


def test_leaf_step_trace_kernel_returns_the_hits_of_the_whole_group_kernel(api, oracle, monkeypatch):
    """The production k_trace_fast tests at most one leaf primitive per step (trace_fast.cuh: LEAF_STEP) and refills below 22 busy
    lanes; variant 3 is the kernel it replaced (whole leaf groups, refill below 20), 2 / 1 other refill thresholds. Per ray the
    candidates and their order are the same, so every hit record must be IDENTICAL: primitive, t, u, v bit for bit — on primary rays,
    rays aimed at the mesh and scattered rays of the headline scene, and on a scene whose BVH holds spheres and boxes."""
    for args, n in ((DRAGON, 300_000), (["-w=600"], 150_000)):
        scene, cam, prm = _load(api, args)
        g = api.GpuScene(scene)
        o = oracle.OracleScene(scene)
        rays, _ = primary_rays(oracle, cam, n, seed=21)
        aimed = _rays_at_mesh(o, rays, n, seed=22) if args[0] == "DRAGON" else np.zeros((0, 6))
        xi = np.random.default_rng(23).uniform(0.02, 0.98, (n, 12))
        b = o.bounce(rays, xi, 0.25)
        cont = np.isin(b["kind"], [1, 2]) & np.isfinite(b["next_dir"]).all(axis=1) & (np.linalg.norm(b["next_dir"], axis=1) > 1e-9)
        secondary = np.concatenate([b["next_origin"][cont], b["next_dir"][cont]], axis=1)
        rr = np.concatenate([rays, aimed, secondary], axis=0)
        hits = {}
        for variant in (4, 3, 2, 1):
            monkeypatch.setenv("RTGPU_TRACE_VARIANT", str(variant))
            hits[variant] = g.trace_fast(rr)
        monkeypatch.delenv("RTGPU_TRACE_VARIANT")
        assert (hits[4]["instance"] >= 0).mean() > 0.5
        for variant in (3, 2, 1):
            for field in ("instance", "primitive", "t", "u", "v"):
                assert np.array_equal(hits[4][field], hits[variant][field]), (args[0], variant, field)


def test_device_render_and_resolve_add(api):
    """rtgpu_render_device (caller-owned fp32 accumulator on a caller stream) + rtgpu_resolve_device, which ADDS the
    resolved f64 Buffer into the host buffer (camera.rs:247-253), agree with rtgpu_render."""
    import torch
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=64", "-s=64", "-t=4"])
    g = api.GpuScene(scene)
    full = g.render(cam, prm)
    spp = prm.replicas * prm.sqrt_spt ** 2
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        accum = torch.zeros((cam.height, cam.width, 4), dtype=torch.float32, device="cuda")
        g.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream)
        out = np.zeros_like(full)
        api.resolve_device(accum.data_ptr(), cam.width, cam.height, spp, 0, stream.cuda_stream, out=out)
        api.resolve_device(accum.data_ptr(), cam.width, cam.height, spp, 0, stream.cuda_stream, out=out)
    assert np.allclose(out, 2.0 * full, rtol=1e-4, atol=1e-5)


def test_gpu_build_handles_duplicate_keys_and_tiny_inputs(api, oracle, monkeypatch, tmp_path):
    """The radix tree tells equal Morton codes apart by their sorted position (Karras 2012, section 4): 60 spheres of which 40
    share ONE centre (identical boxes, identical keys), plus well-separated ones; and inputs of 2 and 3 primitives (a
    radix tree needs at least two leaves; more than 16 analytic primitives so that none stays on the broadcast list).
    First hits must equal the oracle's list scan (ties between coincident spheres go to the first in list order, Q13)."""
    monkeypatch.setenv("RTGPU_BUILD_GPU_MIN", "0")
    monkeypatch.setenv("RTGPU_BROADCAST_MAX", "0")          # everything into the BVH, however few primitives
    rng = np.random.default_rng(3)
    for n_dup, n_other in ((40, 20), (2, 0), (1, 2), (0, 17)):
        lines = ["@config output_width = 96", "@config aspect_ratio = 1", "@config camera_pos = 0,0,12", "@config camera_target = 0,0,0",
                 "m: lambertian (constant 0.7,0.7,0.7)"]
        names = []
        for i in range(n_dup):
            lines.append(f"d{i}: sphere 0.5,0.25,0 0.8 $m"); names.append(f"$d{i}")
        for i in range(n_other):
            c = rng.uniform(-4, 4, 3)
            lines.append(f"s{i}: sphere {c[0]:.3f},{c[1]:.3f},{c[2]:.3f} {rng.uniform(0.2, 0.9):.3f} $m"); names.append(f"$s{i}")
        lines += ["sky: sky (constant 1,1,1)", "world: list " + " ".join(names) + " $sky", "lights: list $sky"]
        (tmp_path / "dups").write_text("\n".join(lines) + "\n")
        scene, cam, prm = api.load([str(tmp_path / "dups"), "-s=16", "-t=1"])
        g = api.GpuScene(scene)
        st = g.stats()
        assert st.bvh_prims == n_dup + n_other and (st.ms_build_device > 0 or st.bvh_prims < 2)
        o = oracle.OracleScene(scene)
        rays, _ = primary_rays(oracle, cam, 30_000, seed=5)
        ho, hg = o.trace_primary(rays), g.trace_primary(rays)
        assert np.array_equal(ho["object"], hg["object"]) and np.array_equal(ho["instance"], hg["instance"]), (n_dup, n_other)
        hit = (ho["object"] >= 0) & (ho["t"] < 1e300)
        assert hit.sum() > 100 and np.array_equal(ho["t"][hit], hg["t"][hit])
        img = g.render(cam, prm)
        assert np.isfinite(img).all() and img[..., :3].mean() > 0.1


@pytest.mark.parametrize("args,n", [(["SCENE:light_test"], 200_000), (DRAGON, 200_000), (["-w=600"], 100_000), (["SCENE:cornell"], 50_000), (["MODEL"], 100_000)],
                         ids=lambda a: a[0] if isinstance(a, list) else None)
def test_gpu_built_acceleration_structure_gives_the_oracles_first_hits(api, oracle, monkeypatch, args, n):
    """rtgpu_scene_create builds the wide BVH on the device (build_gpu.cuh: Morton sort, Karras radix
    tree, bottom-up refit + SAH-optimal collapse tables, level-wise collapse into the same 80-byte quantised nodes). A
    different tree, the same closest hits: ids bit-exact and t bit-equal against the oracle's octree walk in exact mode,
    the production kernel agrees with the exact one as on the host-built tree, and a render is the host-built tree's image."""
    scene, cam, prm = _load(api, args + ["-s=16", "-t=4"] + (["-w=256"] if args[0] != "-w=600" else []))
    monkeypatch.setenv("RTGPU_BUILD", "host")           # the host's binned-SAH builder (bvh.cpp), for comparison
    host = api.GpuScene(scene)
    monkeypatch.delenv("RTGPU_BUILD")
    monkeypatch.setenv("RTGPU_BUILD_GPU_MIN", "0")      # (by default scenes below 131 072 primitives keep the cheap host SAH build)
    g = api.GpuScene(scene)                             # built on the device
    st = g.stats()
    assert host.stats().ms_build_device == 0
    assert st.ms_build_device > 0 or st.bvh_nodes == 0   # (cornell: every primitive is on the broadcast list, nothing to build)
    assert st.bvh_prims == host.stats().bvh_prims
    o = oracle.OracleScene(scene)
    rays, _ = primary_rays(oracle, cam, n, seed=21)
    if args[0] == "DRAGON":
        rays = np.concatenate([rays, _rays_at_mesh(o, rays, n, seed=22)])
    ho, hg = o.trace_primary(rays), g.trace_primary(rays)
    assert np.array_equal(ho["object"], hg["object"]) and np.array_equal(ho["instance"], hg["instance"])
    assert np.array_equal(ho["primitive"], hg["primitive"])
    hit = ho["object"] >= 0
    assert np.array_equal(ho["t"][hit], hg["t"][hit])
    rr = rays.copy()
    rr[:, 3:6] = rr[:, 3:6].astype(np.float32)
    he, hf = g.trace_primary(rr), g.trace_fast(rr)
    surf = (he["object"] >= 0) & np.isfinite(he["t"]) & (he["t"] < 1e300)
    same = (np.where(surf, he["instance"], -1) == hf["instance"]) & ((he["primitive"] == hf["primitive"]) | ~surf)
    assert same.mean() > 0.9997, same.mean()
    a, b = host.render(cam, prm), g.render(cam, prm)
    assert np.abs(a - b).mean() <= 2e-3 * a.mean()        # rays within fp32 rounding of an edge may pick the neighbour
    _record("gpu_build", args[0], {"prims": int(st.bvh_prims), "wide_nodes_gpu": int(st.bvh_nodes), "wide_nodes_host": int(host.stats().bvh_nodes),
                                   "device_build_ms": float(st.ms_build_device), "host_flatten_and_build_ms": float(host.stats().ms_build),
                                   "gpu_flatten_and_build_ms": float(st.ms_build), "fast_vs_exact_id_agreement": float(same.mean())})


def test_fixed_flag_makes_the_estimator_independent_of_the_light_bias(api):
    """`--fixed` (RTGPU_FLAG_FIXED): Plane::random covers the whole quad and Sphere::random uses
    an orthonormal basis, so what is sampled is what pdf_value describes and the mixture estimator is unbiased — the
    converged image must then NOT depend on --light-bias (bias 0 never samples the lights and is unbiased in either mode).
    In the default mode the reference's quarter-area plane sampler (Q1) over-weights one quadrant of the light by up to 4x
    and the image brightens with the bias: parity is defined against the reference, so that stays the default."""
    from rust_raytracer_b200 import _abi as abi
    img, lum = {}, {}
    for fixed in (False, True):
        for bias in ("0", "0.5"):
            scene, cam, prm = _load(api, ["SCENE:cornell", "-w=96", "-s=1024", "-t=4", f"--light-bias={bias}"] + (["--fixed"] if fixed else []))
            assert bool(prm.flags & abi.FLAG_FIXED) == fixed
            im = api.GpuScene(scene).render(cam, prm)
            assert np.isfinite(im).all()
            img[(fixed, bias)] = luminance(im).reshape(12, 8, 12, 8).mean(axis=(1, 3))   # 8 x 8 block means
            lum[(fixed, bias)] = float(luminance(im).mean())
    # the quirk barely moves the MEAN (one quadrant of the light over-weighted 4x, three never sampled: it nearly cancels);
    # its bias is structural — where the light's penumbrae fall — so compare block means between bias 0 and bias 0.5
    d_ref = float(np.sqrt(((img[(False, "0.5")] - img[(False, "0")]) ** 2).mean()))
    d_fix = float(np.sqrt(((img[(True, "0.5")] - img[(True, "0")]) ** 2).mean()))
    print(lum, "block rmse bias 0.5 vs 0: reference", d_ref, "fixed", d_fix)
    _record("fixed_flag", "cornell 96x96 @1024 spp", {"mean_lum": {f"{'fixed' if f else 'reference'} bias {b}": v for (f, b), v in lum.items()},
                                                      "block_rmse_bias_0.5_vs_0": {"reference": d_ref, "fixed": d_fix}})
    assert abs(lum[(False, "0")] - lum[(True, "0")]) <= 1e-6 * lum[(True, "0")]                    # bias 0: the samplers are never used
    assert abs(lum[(True, "0.5")] - lum[(True, "0")]) <= 0.01 * lum[(True, "0")]                   # fixed: independent of the bias
    assert d_ref >= 2.0 * d_fix                                                                    # reference quirks: it is not


def test_render_srgb8_is_render_plus_tonemap_on_the_device(api):
    """rtgpu_render_srgb8 (accumulator -> / spp -> ACES -> sRGB -> u8 on the device, 3 bytes per pixel back) gives the
    pixels of rtgpu_render + rtgpu_tonemap_srgb8 (main.rs:74-82 through the f64 Buffer)."""
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=96", "-s=64", "-t=4"])
    g = api.GpuScene(scene)
    for mode in ("aces", "clamp"):
        a = api.tonemap_srgb8(g.render(cam, prm), mode)
        b = g.render_srgb8(cam, prm, mode)
        assert a.shape == b.shape == (cam.height, cam.width, 3) and b.max() > 100
        d = np.abs(a.astype(int) - b.astype(int))   # fp32 atomics land in a different order run to run
        assert d.max() <= 2 and (d > 0).mean() < 0.05


CLOSED_ROOM = """
@config output_width = 128
@config aspect_ratio = 1
@config camera_pos = 277.5,277.5,100
@config camera_target = 277.5,277.5,555
@config focal_length = 20
white: lambertian (constant 0.9,0.9,0.9)
mirror: metal (constant 0.95,0.95,0.95) (constant 0.0)
floor: plane 277.5,0,277.5 277.5,0,0 0,0,-277.5 $white
ceiling: plane 277.5,555,277.5 277.5,0,0 0,0,277.5 $white
back_wall: plane 277.5,277.5,555 0,277.5,0 277.5,0,0 $mirror
front_wall: plane 277.5,277.5,0 0,277.5,0 -277.5,0,0 $white
left_wall: plane 555,277.5,277.5 0,277.5,0 0,0,-277.5 $white
right_wall: plane 0,277.5,277.5 0,277.5,0 0,0,277.5 $white
light: plane 277.5,554.9,277.5 -30,0,0 0,0,-30 (emissive (constant 40,40,40)) backface
ball: sphere 200,100,300 100 (glass)
world: list $floor $ceiling $back_wall $front_wall $left_wall $right_wall $light $ball
lights: list $light
"""


def test_tail_merging_sizes_from_measured_survival_and_never_writes_out_of_bounds(api, monkeypatch, tmp_path):
    """the arrays that receive the merged tail of a group of batches are no longer assumed large
    enough. A CLOSED room with bright walls, a mirror and --light-bias=0 keeps ~90 % of its paths alive per bounce:
    (1) the pilot sample measures that and the schedule falls back to un-merged tails (stats.tail_group == 1);
    (2) with the survival under-reported on purpose (RTGPU_SURVIVAL_SCALE) the device-side bounds checks trip,
    rtgpu_render notices and repeats the render without merging: same image, no out-of-bounds write;
    (3) rtgpu_render_device (asynchronous, caller's accumulator) reports the overflow from rtgpu_get_stats."""
    (tmp_path / "closed").write_text(CLOSED_ROOM)
    scene, cam, prm = api.load([str(tmp_path / "closed"), "-s=256", "-t=4", "--max-depth=12", "--light-bias=0"])
    monkeypatch.setenv("RTGPU_BATCH_MPATHS", "1")          # 16 384 pixels: 64 samples per batch, 4 batches
    monkeypatch.setenv("RTGPU_LANES", "1")
    monkeypatch.setenv("RTGPU_TAIL_GROUP", "1")
    g = api.GpuScene(scene)
    ref = g.render(cam, prm)
    s = g.stats()
    assert s.rays_per_bounce[5] > 0.5 * s.rays_per_bounce[0]                 # most paths are still alive at bounce 5
    monkeypatch.setenv("RTGPU_TAIL_GROUP", "3")
    g1 = api.GpuScene(scene)
    img = g1.render(cam, prm)
    assert g1.stats().tail_group == 1                                          # (1)
    assert np.abs(img - ref).mean() <= 1e-4 * ref.mean()
    monkeypatch.setenv("RTGPU_SURVIVAL_SCALE", "0.25")                        # (2) pretend a quarter of them survive
    g2 = api.GpuScene(scene)
    img2 = g2.render(cam, prm)
    assert g2.stats().tail_group == 1                                          # the retry ran un-merged
    assert np.isfinite(img2).all() and np.abs(img2 - ref).mean() <= 1e-4 * ref.mean()
    import torch                                                               # (3)
    g3 = api.GpuScene(scene)
    accum = torch.zeros((cam.height, cam.width, 4), dtype=torch.float32, device="cuda")
    g3.render_device(cam, prm, accum.data_ptr(), 0, None)
    with pytest.raises(api.RtgpuError) as e:
        g3.stats()
    assert "path pool overflow" in str(e.value)
    monkeypatch.delenv("RTGPU_SURVIVAL_SCALE")
    # an open, light-dominated scene keeps its merged tails, in far less memory than the worst case
    scene4, cam4, prm4 = _load(api, ["SCENE:light_test", "-w=256", "-s=256", "-t=4"])
    g4 = api.GpuScene(scene4)
    g4.render(cam4, prm4)
    s4 = g4.stats()
    assert s4.tail_group >= 2 and s4.pool_bytes < 368 * s4.batch_paths * s4.lanes


def test_max_depth_beyond_the_statistics_rows(api, oracle):
    """--max-depth is not limited by the 64 statistics rows (config.rs:76 accepts any usize): bounces past 62 are
    summed into the last row."""
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=32", "-s=16", "-t=1", "--max-depth=100"])
    g = api.GpuScene(scene)
    img = g.render(cam, prm)
    p20 = type(prm).from_buffer_copy(prm)
    p20.max_depth = 20
    img20 = g.render(cam, p20)
    assert np.isfinite(img).all() and img[..., :3].mean() >= img20[..., :3].mean() * 0.999


def test_in_process_multi_gpu_render_matches_single(api):
    """rtgpu_render with n_gpus = 2: sample strata sharded over two devices, accumulators summed through
    NVLink peer reads in k_reduce_resolve. Same Philox streams => same image up to fp32 summation order."""
    n = api.lib().rtgpu_device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=96", "-s=64", "-t=4"])
    one = api.GpuScene(scene, devices=[0]).render(cam, prm)
    g2 = api.GpuScene(scene, devices=[0, 1])
    p2 = type(prm).from_buffer_copy(prm)
    p2.n_gpus = 2
    two = g2.render(cam, p2)
    assert np.allclose(one, two, rtol=1e-4, atol=1e-5)
    assert g2.stats().paths == cam.width * cam.height * prm.replicas * prm.sqrt_spt ** 2


def test_image_independent_of_batching_and_lanes(api, monkeypatch):
    """The Philox stream of a path vertex is addressed by (pixel, sample, bounce), so batch size, lane count
    (streams + pools running batches concurrently) and the trace-kernel variant only change the fp32 summation
    order of the accumulator."""
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=160", "-s=64", "-t=4"])   # 25 600 pixels x 64 spp = 1.6 M paths
    g = api.GpuScene(scene)
    monkeypatch.setenv("RTGPU_LANES", "1")
    monkeypatch.setenv("RTGPU_BATCH_MPATHS", "16")
    ref = g.render(cam, prm)
    for lanes, batch_paths, variant in ((4, 1, 4), (3, 1, 2), (1, 1, 0), (4, 16, 1)):
        monkeypatch.setenv("RTGPU_LANES", str(lanes))
        monkeypatch.setenv("RTGPU_BATCH_MPATHS", str(batch_paths))   # 1 Mi paths -> 2 batches of 40 / 24 samples
        monkeypatch.setenv("RTGPU_TRACE_VARIANT", str(variant))
        img = g.render(cam, prm)
        # variant 0 is the v1 kernel (f64 quads): a few edge rays may pick another wall, hence the loose bound there
        tol = 2e-3 if variant == 0 else 1e-4
        assert np.abs(img - ref).mean() <= tol * ref.mean(), (lanes, batch_paths, variant, np.abs(img - ref).mean() / ref.mean())


def test_streamed_records_are_independent_of_grid_and_chunking(api, monkeypatch):
    """Ray records are appended at positions reserved per warp in chunks whose size depends on the grid and on
    the bounce's population; the unused end of a chunk becomes dead records the trace kernel skips
    (kernels.cuh: RecordCursor, routed_flush). Grid sizes therefore change where every record lives and how many
    dead ones exist, and must change neither the image nor any per-bounce ray count."""
    for args in (["SCENE:light_test", "-w=288", "-s=144", "-t=4"], ["SCENE:cornell", "-w=256", "-s=64", "-t=4", "--max-depth=12"]):
        scene, cam, prm = _load(api, args)   # 7.9 M / 4.2 M paths: several 1 Mi-path batches per lane
        g = api.GpuScene(scene)
        runs = []
        # (trace grid, shade grid, lanes, tail-merging group): the batches are 1 Mi paths, so one lane runs several and
        # groups of 2 / 3 of them share their bounces >= RTGPU_TAIL_BOUNCE (api.cu: enqueue_render)
        for trace_ctas, shade_ctas, lanes, tail in ((4, 6, 1, 1), (8, 3, 1, 2), (1, 1, 2, 3), (2, 12, 4, 1), (4, 6, 1, 3)):
            monkeypatch.setenv("RTGPU_TAIL_GROUP", str(tail))
            monkeypatch.setenv("RTGPU_TAIL_BOUNCE", "2" if tail == 2 else "4")
            monkeypatch.setenv("RTGPU_TRACE_CTAS_PER_SM", str(trace_ctas))
            monkeypatch.setenv("RTGPU_SHADE_CTAS_PER_SM", str(shade_ctas))
            monkeypatch.setenv("RTGPU_LANES", str(lanes))
            monkeypatch.setenv("RTGPU_BATCH_MPATHS", "1")
            s0 = g.stats()
            img = g.render(cam, prm)
            s1 = g.stats()
            runs.append((img, [int(s1.rays_per_bounce[i] - s0.rays_per_bounce[i]) for i in range(12)],
                         [int(s1.rays_traversed_per_bounce[i] - s0.rays_traversed_per_bounce[i]) for i in range(12)]))
        ref, rays, trav = runs[0]
        assert rays[0] == cam.width * cam.height * prm.replicas * prm.sqrt_spt ** 2 and sum(trav) <= sum(rays)
        for img, r, t in runs[1:]:
            assert r == rays and t == trav, (args[0], r, rays, t, trav)       # same paths, same routing
            assert np.abs(img - ref).mean() <= 1e-4 * ref.mean(), args[0]      # fp32 summation order only


def test_broadcast_list_and_trace_bypass_leave_the_image_alone(api, monkeypatch):
    """flatten.cpp keeps up to 16 large fp32 quads / spheres out of the BVH; k_raygen / k_shade test them where a ray
    is made, and rays that cannot reach the BVH bounds skip k_trace_fast. Both are pure re-orderings of
    ObjectList::test (list.rs:58-74): switching them off must give the same picture and the same ray counts."""
    for args in (["SCENE:cornell", "-w=128", "-s=64", "-t=4"], ["SCENE:light_test", "-w=128", "-s=16", "-t=4"]):
        scene, cam, prm = _load(api, args)
        runs = {}
        for name, env in (("default", {}), ("no_bypass", {"RTGPU_BYPASS": "0"}), ("no_list", {"RTGPU_BROADCAST_MAX": "0"})):
            for k in ("RTGPU_BYPASS", "RTGPU_BROADCAST_MAX"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            g = api.GpuScene(scene)            # the list is chosen when the scene is flattened
            img = g.render(cam, prm)
            st = g.stats()
            runs[name] = (img, int(st.rays), int(st.rays_traversed), list(st.rays_per_bounce)[:8])
        ref, rays, trav, per_bounce = runs["default"]
        assert trav < rays                                       # the list resolved some rays on its own
        assert runs["no_bypass"][2] == runs["no_bypass"][1]      # without the bypass every ray enters the trace kernel
        assert runs["default"][2] <= runs["no_list"][2] <= runs["no_list"][1]   # no list: only rays that miss the BVH bounds bypass
        for name in ("no_bypass", "no_list"):
            img, r, _, pb = runs[name]
            assert np.abs(img - ref).mean() <= 2e-3 * ref.mean(), (args[0], name)
            assert pb[0] == per_bounce[0] and abs(r - rays) <= 2e-3 * rays, (args[0], name, r, rays)


def test_dropped_misses_and_inline_emission_change_neither_image_nor_ray_counts(api, monkeypatch):
    """Two shortcuts of the wavefront pipeline (kernels.cuh): a ray that misses everything in a scene without sky / sun and with
    a black background is dropped where the miss is found, and a new ray the pre-test resolves onto an Emissive surface with a
    constant emission map is finished by the kernel that makes it. Both only remove a write / queue / read round trip: with
    either switched off the image is the same (fp32 summation order) and every per-bounce ray count is IDENTICAL (the dropped
    rays are counted, RT_CNT_DROPPED). Also with a non-black background, where a miss must still be shaded."""
    for args in (["SCENE:cornell", "-w=128", "-s=64", "-t=4"], ["SCENE:light_test", "-w=160", "-s=64", "-t=4"], DRAGON + ["-w=96", "-s=16", "-t=4"],
                 ["SCENE:cornell", "-w=96", "-s=64", "-t=4", "-b=0.1,0.2,0.4"]):
        scene, cam, prm = _load(api, args)
        runs = []
        for env in ({}, {"RTGPU_DROP_MISSES": "0"}, {"RTGPU_INLINE_EMISSION": "0"}, {"RTGPU_DROP_MISSES": "0", "RTGPU_INLINE_EMISSION": "0"}):
            for k in ("RTGPU_DROP_MISSES", "RTGPU_INLINE_EMISSION"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            g = api.GpuScene(scene)
            img = g.render(cam, prm)
            st = g.stats()
            runs.append((img, [int(st.rays_per_bounce[i]) for i in range(20)], int(st.rays)))
        ref, per_bounce, rays = runs[0]
        assert per_bounce[0] == cam.width * cam.height * prm.replicas * prm.sqrt_spt ** 2
        for img, pb, r in runs[1:]:
            assert pb == per_bounce and r == rays, (args[0], pb[:6], per_bounce[:6])
            assert np.abs(img - ref).mean() <= 1e-4 * ref.mean(), args[0]


def test_edge_cases_small_and_degenerate_inputs(api, oracle, tmp_path):
    # hooks with zero rays
    scene, cam, prm = _load(api, ["SCENE:cornell", "-w=32", "-s=64", "-t=1"])
    g = api.GpuScene(scene)
    assert len(g.trace_primary(np.zeros((0, 6)))) == 0 and len(g.trace_fast(np.zeros((0, 6)))) == 0
    # max-depth 1: only directly visible emitters contribute (ray_color(depth = 0) returns black, camera.rs:283-285)
    p1 = type(prm).from_buffer_copy(prm)
    p1.max_depth = 1
    img = g.render(cam, p1)
    ref, _ = oracle.OracleScene(scene).render(cam, p1, n_threads=0)
    assert np.isfinite(img).all() and img[..., :3].max() > 1.0                      # the light is visible ...
    assert ((img[..., :3].sum(axis=2) > 0) == (ref[..., :3].sum(axis=2) > 0)).mean() > 0.97   # ... in the same pixels
    assert abs(img[..., :3].mean() - ref[..., :3].mean()) <= 0.05 * ref[..., :3].mean()
    # max-depth 0: black image, no kernel may touch an empty bounce queue
    p0 = type(prm).from_buffer_copy(prm)
    p0.max_depth = 0
    assert not g.render(cam, p0).any()
    # a world without any finite primitive (sky only): no BVH at all, every ray is shaded by the miss path
    (tmp_path / "s").write_text("@config output_width = 16\n@config aspect_ratio = 1\nsk: sky (constant 0.5,0.6,0.7)\nworld: list $sk\nlights: list $sk\n")
    s2, c2, q2 = api.load([str(tmp_path / "s"), "-s=4", "-t=1"])
    img2 = api.GpuScene(s2).render(c2, q2)
    assert np.allclose(img2[..., :3], np.array([0.5, 0.6, 0.7]), rtol=1e-5)
    # a 1 x 1 image
    s3, c3, q3 = _load(api, ["SCENE:cornell", "-w=1", "-s=16", "-t=1"])
    img3 = api.GpuScene(s3).render(c3, q3)
    assert img3.shape == (1, 1, 4) and np.isfinite(img3).all()


def test_baseline_config_1_at_full_size_and_spp(api, oracle):
    """BASELINE config 1 — `scenes/cornell` 600x600, `-s=100 -t=10` (10 x 3x3 strata = 90 spp by the README spp rule), max-depth
    20 — is the one config the CPU oracle can afford at FULL size and spp: the same gates as the reduced converged tests, 1:1."""
    args = ["SCENE:cornell", "-w=600", "-s=100", "-t=10", "--max-depth=20"]
    scene, cam, prm = _load(api, args)
    assert (cam.width, cam.height) == (600, 600) and prm.replicas * prm.sqrt_spt ** 2 == 90
    _converged(api, oracle, args)


def _dragon_args(extra):
    return DRAGON + extra


def test_converged_image_parity_headline_scene(api, oracle):
    """BASELINE config 5 (cornell_dragon, ~1.0 M-triangle proxy mesh, glossy material) at reduced size against the
    oracle (octree + f64, the reference's structures): RMSE within 1.5 x the oracle's own noise floor, mean
    luminance within the measured 3-sigma gate."""
    _converged(api, oracle, _dragon_args(["-w=160", "-s=144", "-t=4"]))


def test_full_size_properties_headline_scene(api, oracle, monkeypatch):
    """BASELINE config 5 at FULL size (1200 x 1200, one 20-spp slice of the 1000-spp job = 28.8 M paths, in ten
    2-spp batches on four lanes so that the lanes run tail-merging groups as in the whole job). The oracle cannot
    render this in test time, so check size-independent properties:
    (1) two disjoint sample shards are two independent estimates of the same image (means agree, and their
    difference is pure noise: no structure survives an 8 x 8 box filter beyond the expected sigma);
    (2) shards sum exactly (up to fp32 order) to the render of their union;
    (3) the 8 x 8 box-filtered image agrees with an oracle render of the 150 x 150 image (same camera, so a pixel
    of the small image covers exactly an 8 x 8 block of the large one);
    (4) rays per path lies in the band the scene's albedos imply."""
    monkeypatch.setenv("RTGPU_BATCH_MPATHS", "3")   # 2 samples of 1.44 M pixels per batch
    scene, cam, prm = _load(api, _dragon_args(["-w=1200", "-s=20", "-t=5"]))
    assert (cam.width, cam.height) == (1200, 1200)
    g = api.GpuScene(scene)
    spp = prm.replicas * prm.sqrt_spt ** 2
    assert spp == 20
    halves = []
    for b, e in ((0, spp // 2), (spp // 2, spp)):
        p = type(prm).from_buffer_copy(prm)
        p.sample_begin, p.sample_end = b, e
        halves.append(g.render(cam, p))            # each shard contributes sum / spp (ADD semantics)
    s0 = g.stats()
    full = g.render(cam, prm)
    s1 = g.stats()
    assert np.isfinite(full).all() and (full[..., :3] >= 0).all()
    assert np.allclose(halves[0] + halves[1], full, rtol=2e-4, atol=1e-5)                      # (2)
    a, b = 2.0 * halves[0][..., :3], 2.0 * halves[1][..., :3]
    assert abs(a.mean() - b.mean()) <= 0.01 * full[..., :3].mean()                             # (1) means
    box = lambda im: im.reshape(150, 8, 150, 8, 3).mean(axis=(1, 3))
    d = box(a) - box(b)
    sigma = np.sqrt(((a - b) ** 2).mean() / 64.0)    # noise of a box-filtered difference if pixels are independent
    assert np.sqrt((d ** 2).mean()) <= 1.5 * sigma                                             # (1) no structure
    rays_per_path = (s1.rays - s0.rays) / float(cam.width * cam.height * spp)
    assert 2.8 < rays_per_path < 3.6, rays_per_path                                            # (4)
    small_scene, small_cam, small_prm = _load(api, _dragon_args(["-w=150", "-s=256", "-t=4"]))
    ref, _ = oracle.OracleScene(small_scene).render(small_cam, small_prm, n_threads=0)
    ref2_prm = type(small_prm).from_buffer_copy(small_prm)
    ref2_prm.seed = 777
    ref2, _ = oracle.OracleScene(small_scene).render(small_cam, ref2_prm, n_threads=0)
    floor = rmse(ref, ref2)
    got = rmse(box(full[..., :3]), ref[..., :3])
    lum_g, lum_o = luminance(box(full[..., :3])).mean(), luminance(ref).mean()
    print(f"full-size vs oracle(150x150): rmse {got:.5f}, oracle noise floor {floor:.5f}, lum {lum_g:.5f} / {lum_o:.5f}")
    assert got <= 1.5 * floor + 1e-6                                                           # (3)
    assert abs(lum_g - lum_o) <= 0.02 * lum_o


@pytest.mark.gpu
def test_cli_writes_the_same_png_as_the_api(api, tmp_path):
    """python -m rust_raytracer_b200 <scene> -flags --gpu mirrors main.rs:25-88: progress lines, out.png through the
    Writer::save maths; the file holds exactly the pixels of rtgpu_render + rtgpu_tonemap_srgb8 for the same seed."""
    import subprocess
    import sys
    from PIL import Image
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "out.png"
    flags = [api.scene_path("cornell"), "-w=64", "-s=64", "-t=4", "--max-depth=8"]
    r = subprocess.run([sys.executable, "-m", "rust_raytracer_b200"] + flags + ["--gpu", "--seed=7", f"--out={out}"],
                       cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    assert "Ready:" in r.stdout and "Rendering: 64x64 @64spp on 4 threads (16 samples/thread)" in r.stdout and "Goodbye" in r.stdout
    png = np.asarray(Image.open(out))
    scene, cam, prm = _load(api, flags)
    prm.seed = 7
    img = api.GpuScene(scene).render(cam, prm)
    ref = api.tonemap_srgb8(img, "aces")
    assert png.shape == ref.shape == (64, 64, 3)
    # fp32 atomics accumulate in a different order from run to run: 8-bit values may differ by one step in a few pixels
    d = np.abs(png.astype(int) - ref.astype(int))
    assert d.max() <= 2 and (d > 0).mean() < 0.05


def test_cli_progressive_preview_ends_on_the_same_image(api, tmp_path):
    """--preview=N rewrites the PNG after every N samples per pixel (sample blocks ADDED into one Buffer, rescaled for
    display); the last dump is the image of the single-call render."""
    import subprocess
    import sys
    from PIL import Image
    flags = [api.scene_path("cornell"), "-w=64", "-s=64", "-t=4", "--max-depth=8", "--gpu", "--seed=7"]
    a, b = tmp_path / "a.png", tmp_path / "b.png"
    r = subprocess.run([sys.executable, "-m", "rust_raytracer_b200"] + flags + [f"--out={a}"], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([sys.executable, "-m", "rust_raytracer_b200"] + flags + [f"--out={b}", "--preview=24"], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    assert "24/64 spp" in r.stdout and "48/64 spp" in r.stdout and "64/64 spp" in r.stdout
    d = np.abs(np.asarray(Image.open(a)).astype(int) - np.asarray(Image.open(b)).astype(int))
    assert d.max() <= 2 and (d > 0).mean() < 0.05


def test_native_cli_writes_the_same_png_as_the_python_driver(api, tmp_path):
    """rust_raytracer_b200/rust_raytracer (csrc/cli/main.cpp): main.rs:25-88 as a native program on the C ABI alone. Same
    progress lines and the same out.png as `python -m rust_raytracer_b200`; a scene with a JPEG texture goes through the
    library's own decoder there and through PIL here (texels equal to within the +-1 of two inverse DCTs)."""
    import subprocess
    import sys
    from PIL import Image
    exe = os.path.join(ROOT, "rust_raytracer_b200", "rust_raytracer")
    assert os.path.exists(exe), "run `python __graft_entry__.py` (build) first"
    for scene_name, tol_frac in (("cornell", 0.05), ("earth", 0.08), ("texture_test", 0.05)):
        flags = [api.scene_path(scene_name), "-w=64", "-s=64", "-t=4", "--max-depth=8", "--gpu", "--seed=7"]
        a, b = tmp_path / f"{scene_name}_py.png", tmp_path / f"{scene_name}_native.png"
        r = subprocess.run([sys.executable, "-m", "rust_raytracer_b200"] + flags + [f"--out={a}"], cwd=ROOT, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        n = subprocess.run([exe] + flags + [f"--out={b}"], cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
        assert n.returncode == 0, n.stderr
        for line in ("Ready:", "Rendering: 64x", "@64spp on 4 threads (16 samples/thread) [1 GPU(s)]", "Writing output to file...", "Goodbye :)"):
            assert line in n.stdout, (line, n.stdout)
        pa, pb = np.asarray(Image.open(a)).astype(int), np.asarray(Image.open(b)).astype(int)
        d = np.abs(pa - pb)
        print(f"{scene_name}: native vs python driver, max 8-bit difference {d.max()}, pixels that differ {(d > 0).mean():.4f}")
        assert pa.shape == pb.shape and d.max() <= 3 and (d > 0).mean() < tol_frac
    # progressive dumps and the bare --fixed flag are understood as well
    flags = [api.scene_path("cornell"), "-w=48", "-s=32", "-t=2", "--max-depth=6", "--gpu", "--fixed", "--preview=12", f"--out={tmp_path / 'p.png'}"]
    n = subprocess.run([exe] + flags, cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert n.returncode == 0, n.stderr
    assert "12/32 spp" in n.stdout and "32/32 spp" in n.stdout and os.path.exists(tmp_path / "p.png")


def test_light_test_full_job_matches_the_reference_binarys_output(api):
    """BASELINE config 4 end to end against an OUTPUT OF THE REFERENCE BINARY: samples/sample1.png is what
    `rust_raytracer scenes/light_test -w=2400 -s=1000 -t=10` wrote (README.md:12, ~20 min on an M3 Pro); its 8x8 block
    means are committed in tests/golden/reference_sample1_blocks.npz (make_reference_sample_fixture.py). The scene is
    deterministic, so the two 1000-spp images differ by Monte-Carlo noise only: channel means of the 8-bit image
    within 0.5 %, RMSE of the 8x8 block means <= 0.004 (measured 0.0018 on a 0..1 scale), no block off by more
    than 0.06 (measured 0.024; the largest differences sit on the specular highlights)."""
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_sample1_blocks.npz"))
    scene, cam, prm = _load(api, ["SCENE:light_test", "-w=2400", "-s=1000", "-t=10"])
    assert (cam.width, cam.height) == (2400, 1600) and prm.replicas * prm.sqrt_spt ** 2 == 1000
    img = api.GpuScene(scene).render(cam, prm)
    rgb = api.tonemap_srgb8(img, "aces").astype(np.float64) / 255.0        # Writer::save maths on the GPU
    ref_mean = gold["channel_mean"]
    mean = rgb.mean(axis=(0, 1))
    print("mean sRGB: reference binary", ref_mean, "gpu", mean)
    assert np.all(np.abs(mean - ref_mean) <= 0.005 * ref_mean), (mean, ref_mean)
    blocks = rgb.reshape(200, 8, 300, 8, 3).mean(axis=(1, 3))
    d = blocks - gold["blocks8"].astype(np.float64) / 65535.0
    print("8x8 block rmse", np.sqrt((d ** 2).mean()), "max", np.abs(d).max())
    assert np.sqrt((d ** 2).mean()) <= 0.004
    assert np.abs(d).max() <= 0.06


def test_cornell_dragon_full_job_matches_the_reference_binarys_output_above_the_mesh(api):
    """BASELINE config 5 against an OUTPUT OF THE REFERENCE BINARY: samples/sample2.png is what
    `rust_raytracer scenes/cornell_dragon -w=1200 -s=1000 -t=10 --max-depth=20` wrote (README.md:13). The dragon mesh is
    missing upstream and a proxy is rendered here, so only the picture ABOVE the mesh is compared (top 408 rows: ceiling,
    the area light with its quarter-area sampling quirk plane.rs:120-126, upper walls, frame; 8x8 block means in
    tests/golden/reference_sample2_top_blocks.npz, made by make_reference_sample_fixture.py). The real dragon is a
    larger white body than the proxy and bounces more light back up, so the ceiling is a few per cent brighter in the
    reference; the gates below hold that effect (measured: band mean -3 %, block rmse 0.03) and would not hold a wrong
    light size, wall colour, light-sampling rule or camera (those move the band by 0.1 and more). The whole job is one
    device-resident call (rtgpu_render_srgb8)."""
    gold = np.load(os.path.join(ROOT, "tests", "golden", "reference_sample2_top_blocks.npz"))
    scene, cam, prm = _load(api, DRAGON + ["-w=1200", "-s=1000", "-t=10"])
    assert (cam.width, cam.height) == (1200, 1200) and prm.replicas * prm.sqrt_spt ** 2 == 1000
    rgb = api.GpuScene(scene).render_srgb8(cam, prm, "aces").astype(np.float64) / 255.0
    top = rgb[:408]
    ref = gold["blocks8_top"].astype(np.float64) / 65535.0
    blocks = top.reshape(51, 8, 150, 8, 3).mean(axis=(1, 3))
    d = blocks - ref
    mean, ref_mean = top.mean(axis=(0, 1)), gold["channel_mean_top"]
    stats = {"band_mean_gpu": mean.tolist(), "band_mean_reference": ref_mean.tolist(), "block_rmse": float(np.sqrt((d ** 2).mean())),
             "block_max_abs": float(np.abs(d).max())}
    print(stats)
    _record("reference_images", "sample2.png top band (cornell_dragon, proxy mesh)", stats)
    assert np.all(np.abs(mean - ref_mean) <= 0.06 * ref_mean), (mean, ref_mean)
    assert np.sqrt((d ** 2).mean()) <= 0.04
    assert (np.abs(d) > 0.15).mean() < 0.005
    # the light itself and the frame do not depend on the mesh at all: saturated white / black in both images
    lit = ref.min(axis=2) > 0.97
    assert lit.sum() > 200 and (blocks[lit].min(axis=1) > 0.95).all()
    dark = ref.max(axis=2) < 0.01
    assert dark.sum() > 200 and (blocks[dark].max(axis=1) < 0.02).all()
    # wall colours: left green, right red (the reference's, not mirrored)
    assert blocks[30, 10, 1] > 1.5 * blocks[30, 10, 0] and blocks[30, 140, 0] > 1.5 * blocks[30, 140, 1]
