#!/usr/bin/env python
"""bench.py — headline benchmark of the rtgpu path-tracing backend.

Metric (BASELINE.json): Mpaths/s (and Mrays/s, RMSE) on `scenes/cornell_dragon` at 1200x1200, 1000 spp,
max-depth 20, light-bias 0.25 — the ~870 k-triangle dragon is missing from the reference mount, so the
deterministic ~1.0 M-triangle PROXY mesh of tools/make_dragon_proxy.py is used.

A "step" is one pass of the hot path (Camera::render) over the whole job per GPU: 10 replicas x 10x10 strata =
1000 spp in one render call. With N GPUs the driver-facing `value` is WEAK scaling (every rank renders its own
1000 spp, one NCCL reduce of the fp32 accumulators per step); the same run also measures the north_star job
itself — 1000 spp split over the N ranks — and reports it as `strong`, with a pixel check of the reduced image.

    python bench.py --gpus 1 --steps 5 --warmup 3            # this backend (CUDA)
    python bench.py --config C4                              # another BASELINE config as the timed workload
    python bench.py --impl reference --steps 3 --warmup 1     # CPU oracle on the host cores
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs as concrete runs; b_ray = algorithmic bytes per ray (SURVEY.md §8d:
# 160 state + 80 x ceil(log8 N_prim) node + 48 triangle); published = Mpaths/s derived from the reference README's
# timings on an Apple M3 Pro, None where the reference publishes nothing.
CONFIGS = {
    "C1": {"args": ["SCENE:cornell", "-w=600", "-s=100", "-t=10", "--max-depth=20"], "b_ray": 384, "published": None,
           "label": "scenes/cornell 600x600 @100 spp (-t=10: 10 x 3x3 strata = 90 spp, README spp rule) max-depth 20",
           "slice": ["-w=128", "-s=256", "-t=16"]},
    "C2": {"args": ["-w=1200", "-s=4000", "-t=10", "--light-bias=0.25"], "b_ray": 608, "published": 9.6,
           "label": "default scene (Suzanne + ~440 spheres, sky + sun; scene seed 42) 1200x800 @4000 spp light-bias 0.25",
           "slice": ["-w=128", "-s=144", "-t=16"]},
    "C3a": {"args": ["SCENE:earth", "-w=1200", "-s=1000", "-t=10"], "b_ray": 304, "published": None,
            "label": "scenes/earth 1200x800 @1000 spp", "slice": ["-w=128", "-s=144", "-t=16"]},
    "C3b": {"args": ["SCENE:perlin", "-w=1200", "-s=1000", "-t=10"], "b_ray": 608, "published": None,
            "label": "scenes/perlin 1200x800 @1000 spp", "slice": ["-w=128", "-s=144", "-t=16"]},
    "C3c": {"args": ["SCENE:texture_test", "-w=1200", "-s=1000", "-t=10"], "b_ray": 608, "published": None,
            "label": "scenes/texture_test 1200x800 @1000 spp", "slice": ["-w=128", "-s=144", "-t=16"]},
    "C4": {"args": ["SCENE:light_test", "-w=2400", "-s=1000", "-t=10"], "b_ray": 608, "published": 3.2,
           "label": "scenes/light_test 2400x1600 @1000 spp", "slice": ["-w=160", "-s=256", "-t=16"]},
    "C5": {"args": ["SCENE:cornell_dragon", "-w=1200", "-s=1000", "-t=10", "--max-depth=20", "--light-bias=0.25"], "b_ray": 768,
           "published": 0.585, "label": "scenes/cornell_dragon (PROXY mesh) 1200x1200 @1000 spp max-depth 20 light-bias 0.25",
           "slice": ["-w=160", "-s=144", "-t=16"]},
}
ISSUE_PEAK = 148 * 4 * 1.965e9  # warp instructions / s: 148 SMs x 4 schedulers x 1 per clock x 1.965 GHz


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.samples.append([x.strip() for x in line.split(",")])
        except Exception:  # noqa: BLE001
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0])); mx = max(mx, float(s[1]))
                for nme, v in zip(names, s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:  # noqa: BLE001
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def config_args(api, name, override=()):
    """argv of a BASELINE config; later flags win (config.rs:81-152), so `override` re-sizes the job."""
    cfg = CONFIGS[name]
    if name == "C5":
        from tools.make_dragon_proxy import ensure
        ensure()
    return [api.scene_path(a[6:]) if a.startswith("SCENE:") else a for a in cfg["args"]] + list(override)


def load_workload(api, name, replicas=None, sqrt_spt=None, width=None):
    over = []
    if width:
        over.append(f"-w={width}")
    if replicas and sqrt_spt:
        over += [f"-t={replicas}", f"-s={replicas * sqrt_spt * sqrt_spt}"]
    scene, cam, prm = api.load(config_args(api, name, over))
    if replicas and sqrt_spt:
        assert prm.replicas * prm.sqrt_spt ** 2 == replicas * sqrt_spt * sqrt_spt
    return scene, cam, prm


def cpu_sample(oracle_mod, scene, cam, prm_base, cores, budget_s):
    """Time the oracle on a bounded sample: `cores` replicas x 1 stratum, every row_step-th row."""
    chk = oracle_mod.OracleScene(scene)
    prm = type(prm_base).from_buffer_copy(prm_base)
    prm.replicas, prm.sqrt_spt, prm.sample_begin, prm.sample_end = cores, 1, 0, 0
    # calibrate on a thin slice
    t_cal = 1e30
    for _ in range(2):  # (the first call also pays for first-touch of the per-thread buffers)
        t0 = time.perf_counter()
        _, _ = chk.render_rows(cam, prm, 0, 64, n_threads=cores)
        t_cal = min(t_cal, time.perf_counter() - t0)
    rows_cal = len(range(0, cam.height, 64))
    per_row = t_cal / rows_cal
    rows = max(rows_cal, min(cam.height, int(budget_s / per_row)))
    row_step = max(1, cam.height // rows)
    if row_step == 1:
        # the whole image fits the budget: time it once and add replicas per pixel (each host thread then renders several,
        # camera.rs:197-241) until the sample fills the budget
        t0 = time.perf_counter()
        chk.render_rows(cam, prm, 0, 1, n_threads=cores)
        full = time.perf_counter() - t0
        prm.replicas = cores * max(1, min(8, int(budget_s / max(full, 1e-3))))
    return chk, prm, row_step


def run_reference(args):
    """--impl reference: the CPU oracle (port of the reference; Rust is not buildable here) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from rust_raytracer_b200 import api
    import oracle
    cores = os.cpu_count() or 1
    cfg = CONFIGS[args.config]
    scene, cam, prm = load_workload(api, args.config, 10, 2, args.width)
    budget = max(2.0, min(15.0, 150.0 / (args.steps + args.warmup)))
    chk, sp, row_step = cpu_sample(oracle, scene, cam, prm, cores, budget)
    rows = len(range(0, cam.height, row_step))
    paths = rows * cam.width * sp.replicas
    for _ in range(args.warmup):
        chk.render_rows(cam, sp, 0, row_step, n_threads=cores)
    t0 = time.perf_counter()
    rays = 0
    for i in range(args.steps):
        _, r = chk.render_rows(cam, sp, i % row_step, row_step, n_threads=cores)
        rays += r
    dt = time.perf_counter() - t0
    value = paths * args.steps / dt / 1e6
    sample = f"{sp.replicas} replicas x 1 stratum on every {row_step}-th row of {cam.width}x{cam.height} ({paths} paths/step, {cores} threads)"
    out = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": value / cfg["published"] if cfg["published"] else None, "dtype": "f64",
           "data": "synthetic (proxy dragon mesh)" if args.config == "C5" else "synthetic (reference scene files, seeded RNG)",
           "config": {"workload": cfg["label"], "name": args.config, "sample": sample},
           "mrays_per_s": rays / dt / 1e6,
           "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def image_parity(a, b):
    """rmse and relative mean-luminance difference of two (H, W, >=3) linear images."""
    import numpy as np
    lum = lambda im: 0.2126 * im[..., 0] + 0.7152 * im[..., 1] + 0.0722 * im[..., 2]
    return {"rmse": float(np.sqrt(np.mean((a[..., :3] - b[..., :3]) ** 2))),
            "mean_lum_rel": float(abs(lum(a).mean() - lum(b).mean()) / max(lum(b).mean(), 1e-30))}


def oracle_slice_parity(api, oracle, name, cores):
    """RMSE / mean luminance of the GPU path against the CPU oracle on a reduced slice of config `name` (same scene,
    camera and flags, smaller image and spp so that the oracle finishes in seconds), next to the oracle's own noise floor
    (two seeds). The RNG streams differ by construction, so this is statistical."""
    cfg = CONFIGS[name]
    scene, cam, prm = api.load(config_args(api, name, cfg["slice"]))
    chk = oracle.OracleScene(scene)
    a, _ = chk.render(cam, prm, n_threads=cores)
    p2 = type(prm).from_buffer_copy(prm)
    p2.seed = 12345
    b, _ = chk.render(cam, p2, n_threads=cores)
    gpu = api.GpuScene(scene)
    g = gpu.render(cam, prm)
    gpu.close()
    out = image_parity(g, a)
    out["oracle_noise_floor_rmse"] = image_parity(b, a)["rmse"]
    out["oracle_seed_to_seed_mean_lum_rel"] = image_parity(b, a)["mean_lum_rel"]
    out["slice"] = f"{cam.width}x{cam.height} @{prm.replicas * prm.sqrt_spt ** 2} spp, same scene / camera / depth / bias"
    return out


def time_config(api, torch, name, stream):
    """One BASELINE config at FULL size on this GPU: device-timed whole-job renders."""
    cfg = CONFIGS[name]
    t0 = time.perf_counter()
    scene, cam, prm = api.load(config_args(api, name))
    gpu = api.GpuScene(scene)
    ready = time.perf_counter() - t0
    W, H, spp = cam.width, cam.height, prm.replicas * prm.sqrt_spt ** 2
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    gpu.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream)   # warm-up: pool allocation, pilot sample
    torch.cuda.synchronize()
    s0 = gpu.stats()
    reps = 2
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        accum.zero_()
        gpu.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream)
    e1.record(stream)
    torch.cuda.synchronize()
    s1 = gpu.stats()
    ms = e0.elapsed_time(e1) / reps
    paths = W * H * spp
    rec = {"workload": cfg["label"], "paths_per_job": paths, "ms_per_job": ms, "mpaths_per_s": paths / ms / 1e3,
           "mrays_per_s": (s1.rays - s0.rays) / reps / ms / 1e3, "rays_per_path": (s1.rays - s0.rays) / reps / paths,
           "ready_s": ready, "pool_gb": s1.pool_bytes / 2 ** 30, "tail_group": s1.tail_group,
           "reference_published_mpaths_per_s": cfg["published"]}
    del accum
    gpu.close()
    return rec


_STDOUT_FD = None


def quiet_stdout():
    """Libraries print to fd 1 (NCCL its version line): the contract is ONE JSON line on stdout, so fd 1 points to stderr
    until emit() restores it."""
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(out):
    global _STDOUT_FD
    sys.stdout.flush()
    if _STDOUT_FD is not None:
        os.dup2(_STDOUT_FD, 1)
        os.close(_STDOUT_FD)
        _STDOUT_FD = None
    print(json.dumps(out), flush=True)


def run_ours(args):
    quiet_stdout()
    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    from rust_raytracer_b200 import api, sharding, _abi as abi

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this backend has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = CONFIGS[args.config]
    T, k = 10, args.sqrt_spt
    if args.config == "C1" and args.sqrt_spt == 10:
        k = 3  # -t=10 -s=100
    if args.config == "C2" and args.sqrt_spt == 10:
        k = 20  # -t=10 -s=4000
    spp_rank = T * k * k
    if rank == 0 and args.config == "C5":
        from tools.make_dragon_proxy import ensure
        ensure()
    if world > 1:
        dist.barrier()
    t_load = time.perf_counter()
    scene, cam, prm_all = load_workload(api, args.config, T * world, k, args.width)
    t_load = time.perf_counter() - t_load
    prm = sharding.shard_params(prm_all, rank, world)   # WEAK: this rank's block of the (T*world)*k*k sample strata
    assert prm.sample_end - prm.sample_begin == spp_rank
    t_create = time.perf_counter()
    gpu = api.GpuScene(scene, devices=[local])
    t_create = time.perf_counter() - t_create
    W, H = cam.width, cam.height
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    paths_step_rank = W * H * spp_rank
    # a dedicated stream: the null handle of torch's default stream would make the library use its own
    # stream, and the CUDA events below would not bracket the kernels
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)

    def step():
        accum.zero_()
        gpu.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)  # sum of the per-GPU fp32 accumulators (camera.rs:243-255)

    for _ in range(max(args.warmup, 0)):
        step()
    torch.cuda.synchronize()
    s0 = gpu.stats()
    rays0, launches0 = s0.rays, s0.kernel_launches
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start(); time.sleep(0.3)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    s1 = gpu.stats()
    rays_t = torch.tensor([float(s1.rays - rays0), float(s1.kernel_launches - launches0)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays_t, op=dist.ReduceOp.SUM)
    ms = float(t.item())
    clocks = sampler.finish() if rank == 0 else None
    rays_total, launches = [float(x) for x in rays_t.tolist()]
    value = paths_step_rank * world * args.steps / (ms / 1e3) / 1e6
    # rank 0 rendered global samples [0, spp_rank): exactly the samples of the one-GPU job (same Philox streams)
    weak_rank0 = None
    if world > 1:
        accum.zero_()
        gpu.render_device(cam, prm, accum.data_ptr(), 0, stream.cuda_stream)
        torch.cuda.synchronize()
        if rank == 0:
            weak_rank0 = (accum[..., :3].double() / spp_rank).cpu().numpy()

    # ---- end to end through the host-buffer C ABI call (Camera::render drop-in) -----------------
    total_spp = spp_rank * world
    host = np.zeros((H, W, 4), dtype=np.float64)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()

    def e2e_step():
        if world == 1:
            gpu.render(cam, prm, out=host)          # rtgpu_render: kernels + D2H of the fp32 sums, host scale + add into the f64 Buffer
        else:
            step()
            if rank == 0:
                api.resolve_device(accum.data_ptr(), W, H, total_spp, 0, stream.cuda_stream, out=host)
            torch.cuda.synchronize()

    e2e_step()                                      # untimed: first touch of the host buffer, pinned staging allocation
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_steps = max(1, args.steps)
    e2e_each = []
    for _ in range(e2e_steps):
        ts = time.perf_counter()
        e2e_step()
        e2e_each.append(round((time.perf_counter() - ts) * 1e3, 2))
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_value = paths_step_rank * world * e2e_steps / float(tt.item()) / 1e6

    # ---- STRONG scaling: the north_star job itself. Total fixed at T*k*k spp; rank r renders the contiguous block
    # [r*S/N, (r+1)*S/N) of the samples (camera.rs:197-255 is the replica split being sharded), one NCCL reduce, rank 0
    # resolves into the host f64 Buffer. Timed per job from a barrier to the resolved buffer on rank 0.
    prm_job = type(prm_all).from_buffer_copy(prm_all)
    prm_job.replicas = T                                           # S = T*k*k samples in total, not per rank
    prm_strong = sharding.shard_params(prm_job, rank, world)
    S = T * k * k
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    strong_jobs, first_job_ms = [], None
    host_job = np.zeros((H, W, 4), dtype=np.float64)
    for j in range(2 + max(2, args.steps)):
        host_job[:] = 0.0                          # camera.create_buffer (main.rs:74), outside the timed job like the reference's
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ev[0].record(stream)
        accum.zero_()
        gpu.render_device(cam, prm_strong, accum.data_ptr(), 0, stream.cuda_stream)
        ev[1].record(stream)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        ev[2].record(stream)
        t_res = 0.0
        if rank == 0:
            stream.synchronize()
            tr = time.perf_counter()
            api.resolve_device(accum.data_ptr(), W, H, S, 0, stream.cuda_stream, out=host_job)
            t_res = (time.perf_counter() - tr) * 1e3
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        rec = torch.tensor([wall, ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), t_res], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(rec, op=dist.ReduceOp.MAX)
        if j == 0:
            first_job_ms = float(rec[0])
        elif j >= 2:
            strong_jobs.append([float(x) for x in rec.tolist()])
    med = lambda i: statistics.median(r[i] for r in strong_jobs)
    strong = {"total_spp": S, "spp_per_gpu": prm_strong.sample_end - prm_strong.sample_begin, "jobs_timed": len(strong_jobs),
              "job_ms": med(0), "mpaths_per_s": W * H * S / med(0) / 1e3,
              "render_ms": med(1), "nccl_reduce_ms": med(2) if world > 1 else 0.0, "d2h_resolve_ms": med(3),
              "first_job_ms_incl_allocation": first_job_ms,
              "t1_ms_same_run": ms / args.steps,
              "efficiency": (ms / args.steps) / (world * med(0)),
              "note": "job = barrier .. resolved f64 Buffer on rank 0 (wall clock, max over ranks); render_ms / nccl_reduce_ms = CUDA events "
                      "on the launch stream, max over ranks; t1_ms_same_run = device time of ONE rank rendering all S samples in this run "
                      "(the weak step; at N > 1 it includes the weak NCCL reduce); efficiency = t1 / (N x job)"}
    parity = None
    if world > 1 and rank == 0:
        job_img = host_job[..., :3]
        parity = image_parity(job_img, weak_rank0)
        parity["note"] = (f"{world}-rank job image (NCCL-reduced, resolved) vs the same {S} samples rendered by rank 0 alone: "
                          "same Philox streams, fp32 summation order differs")
        parity["rmse_rel_to_mean"] = parity["rmse"] / float(weak_rank0.mean())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- per-kernel times: one extra step with RTGPU_FLAG_PROFILE, which records a CUDA event pair
    # around every launch ON ITS LAUNCH STREAM and runs the batches on ONE lane, so that each
    # k_trace_fast launch is timed alone (in the timed region above four batches overlap on four
    # streams, and per-launch event spans would include the other lanes' kernels)
    pprof = type(prm).from_buffer_copy(prm)
    pprof.flags |= abi.FLAG_PROFILE
    sp0 = gpu.stats()
    accum.zero_()
    gpu.render_device(cam, pprof, accum.data_ptr(), 0, stream.cuda_stream)
    torch.cuda.synchronize()
    sp1 = gpu.stats()
    # rays k_trace_fast walked through the BVH; the rest of the step's rays were resolved completely by the
    # broadcast pre-test in k_raygen / k_shade and never entered the kernel (they count in `value`, not here)
    prof_rays = float(sp1.rays_traversed - sp0.rays_traversed)
    prof_rays_all = float(sp1.rays - sp0.rays)
    nb = min(pprof.max_depth, 64)
    rays_per_bounce = [int(sp1.rays_per_bounce[i] - sp0.rays_per_bounce[i]) for i in range(nb)]
    traversed_per_bounce = [int(sp1.rays_traversed_per_bounce[i] - sp0.rays_traversed_per_bounce[i]) for i in range(nb)]
    ms_trace, ms_shade, ms_gen = sp1.ms_trace - sp0.ms_trace, sp1.ms_shade - sp0.ms_shade, sp1.ms_raygen - sp0.ms_raygen
    trace_launches = int(sp1.trace_launches - sp0.trace_launches)   # counted by the library where it launches them
    peak, peak_src = load_peaks()
    b_ray = cfg["b_ray"]
    achieved = prof_rays * b_ray / (ms_trace / 1e3) / 1e9 if ms_trace > 0 else None
    traffic = None
    issue = {}
    try:  # ncu --set full capture of one k_trace_fast launch at the production grid (tools/ncu_summary.py writes the file)
        with open(os.path.join(ROOT, "profiles", "trace_traffic.json")) as f:
            cap = json.load(f)
        traffic = cap.get("dram_bytes_per_ray") * prof_rays / trace_launches
        if cap.get("inst_executed") and cap.get("duration_ms"):
            issue = {"issue_frac": cap["inst_executed"] / (cap["duration_ms"] / 1e3 * ISSUE_PEAK), "active_lanes": cap.get("active_lanes"),
                     "warp_inst_per_ray": cap["inst_executed"] / cap["rays_in_launch"], "issue_capture": cap.get("capture")}
    except Exception:  # noqa: BLE001
        pass
    st = gpu.stats()
    roofline = {"bound": "hbm", "kernel": "k_trace_fast", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_ray": b_ray, "launches": trace_launches,
                "rays_per_launch": prof_rays / max(trace_launches, 1),
                "note": "per launch (mean over the %d k_trace_fast launches of one profiled step): achieved = rays the kernel traversed x %d B / "
                        "CUDA-event duration on the launch stream, launches timed alone; traffic = ncu dram bytes per ray "
                        "(profiles/trace_traffic.json) x rays per launch, in bytes; issue_frac = smsp__inst_executed / (t x 1.163e12 warp "
                        "inst/s), active_lanes = smsp__thread_inst_executed_per_inst_executed of the same capture (SURVEY.md 8d: the issue "
                        "roofline is the co-bound of this kernel)" % (trace_launches, b_ray)}
    roofline.update(issue)
    out = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": value / cfg["published"] if cfg["published"] else None, "dtype": "f32",
        "vs_baseline_note": "BASELINE.md section 1: Mpaths/s derived from the reference README's timing of this config on an "
                            "Apple M3 Pro (10 threads)" + ("; the reference used the real dragon mesh, here a proxy mesh of the same size, see `data`"
                                                           if args.config == "C5" else ""),
        "data": "synthetic (proxy dragon mesh; the reference's dragon_high.obj is missing upstream)" if args.config == "C5"
                else "synthetic (reference scene files; scene RNG seed 42)",
        "config": {"workload": cfg["label"] + f" [{st.bvh_triangles} tris], step = {spp_rank} spp per GPU in one render call "
                               f"({T} replicas x {k}x{k} strata)", "name": args.config,
                   "paths_per_step": paths_step_rank * world,
                   "l2": f"working set (path pools {st.pool_bytes / 2 ** 30:.1f} GB: streamed ray records; scene ~200 MB) exceeds the 126 MB L2",
                   "parallelism": f"sample strata sharded over {world} GPU(s), one NCCL reduce of the fp32 accumulator per step"
                   if world > 1 else "1 GPU",
                   "schedule": {"lanes": st.lanes, "batch_paths": st.batch_paths, "tail_group": st.tail_group, "tail_bounce": st.tail_bounce,
                                "pool_gb": st.pool_bytes / 2 ** 30}},
        "mrays_per_s": rays_total / (ms / 1e3) / 1e6, "rays_per_path": rays_total / (paths_step_rank * world * args.steps),
        "kernel_ms_one_step_profiled": {"trace": ms_trace, "shade": ms_shade, "raygen": ms_gen,
                                        "note": "rank 0, one extra step on ONE lane with an event pair per launch; the timed "
                                                "region overlaps 4 batches on 4 streams and is shorter than this sum"},
        "rays_per_bounce_one_step": rays_per_bounce,
        "rays_traversed_per_bounce_one_step": traversed_per_bounce,
        "rays_resolved_by_broadcast_pretest": 1.0 - prof_rays / prof_rays_all if prof_rays_all > 0 else None,
        "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": "Mpaths/s", "ms_each_step_rank0": e2e_each, "h2d_bytes_per_step": C.sizeof(abi.CameraDesc) + C.sizeof(abi.RenderParams),
                "d2h_bytes_per_step": W * H * 4 * 4 if world == 1 else W * H * 4 * 8,
                "note": "rtgpu_render(host f64 Buffer): scene uploaded once beforehand like the reference's 'Ready' phase; one GPU copies the "
                        "fp32 sums (16 B / pixel) and scales + adds on the host, N GPUs resolve to f64 on rank 0 and copy 32 B / pixel"},
        "strong": strong, "parity": parity,
        "gpu_launches": int(launches), "clocks": clocks,
        "build": {"bvh_nodes": st.bvh_nodes, "bvh_prims": st.bvh_prims, "ms_flatten_and_bvh": st.ms_build,
                  "ms_scene_dsl_and_obj_parse": t_load * 1e3, "ms_scene_create_incl_upload": t_create * 1e3,
                  "note": "the reference's 'Ready' phase, outside the timed region; multi-threaded host code"},
    }
    if world == 1 and not args.no_cpu:
        import oracle
        cores = os.cpu_count() or 1
        del accum
        gpu.close()                                  # the other configs need the memory of this scene's pools
        chk, sp, row_step = cpu_sample(oracle, scene, cam, prm, cores, 12.0)
        rows = len(range(0, H, row_step))
        t0 = time.perf_counter()
        _, r = chk.render_rows(cam, sp, 0, row_step, n_threads=cores)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": rows * W * sp.replicas / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                               "mrays_per_s": r / dt / 1e6,
                               "sample": f"{sp.replicas} replicas x 1 stratum on every {row_step}-th row of {W}x{H} ({rows * W * sp.replicas} paths, {dt:.1f} s, {cores} threads)"}
        del chk
        # the metric names RMSE: GPU vs oracle on a reduced slice of the timed config
        out["rmse_vs_oracle"] = oracle_slice_parity(api, oracle, args.config, cores)
        # every BASELINE config at full size on this GPU (whole-job device time), each with its oracle slice
        if not args.no_configs:
            table = {}
            for name in CONFIGS:
                rec = {"workload": CONFIGS[name]["label"]}
                if name == args.config and args.width is None and spp_rank == T * k * k:
                    rec.update({"paths_per_job": paths_step_rank, "ms_per_job": ms / args.steps, "mpaths_per_s": value,
                                "mrays_per_s": out["mrays_per_s"], "rays_per_path": out["rays_per_path"],
                                "reference_published_mpaths_per_s": cfg["published"], "note": "the timed workload of this line"})
                    rec["parity_vs_oracle_slice"] = out["rmse_vs_oracle"]
                else:
                    rec = time_config(api, torch, name, stream)
                    rec["parity_vs_oracle_slice"] = oracle_slice_parity(api, oracle, name, cores)
                table[name] = rec
            out["configs"] = table
    emit(out)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C5", choices=sorted(CONFIGS), help="BASELINE config that is the timed workload (default C5, the headline)")
    ap.add_argument("--width", type=int, default=None, help="override the config's image width")
    ap.add_argument("--sqrt-spt", type=int, default=10, help="strata per replica = sqrt_spt^2; spp per GPU per step = 10 * sqrt_spt^2 "
                    "(default 10: one step = the whole BASELINE job, 10 replicas x 10x10 strata = 1000 spp, in one render call per GPU)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg, the oracle slices and the configs table")
    ap.add_argument("--no-configs", action="store_true", help="skip the table of the other BASELINE configs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
