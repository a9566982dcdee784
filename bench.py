#!/usr/bin/env python
"""bench.py — headline benchmark of the rtgpu path-tracing backend.

Metric (BASELINE.json): Mpaths/s (and Mrays/s) on `scenes/cornell_dragon` at 1200x1200,
max-depth 20, light-bias 0.25 — the ~870 k-triangle dragon is missing from the reference
mount, so the deterministic ~1.0 M-triangle PROXY mesh of tools/make_dragon_proxy.py is used.

A "step" is one pass of the hot path (Camera::render) over `--spp` samples per pixel per GPU
(default 40 = 10 replicas x 2x2 strata, i.e. 57.6 M camera paths per GPU per step); throughput
in paths/s does not depend on the slice size because cost is linear in spp (camera.rs:213-228).
With N GPUs every rank renders its own disjoint slice of the sample strata (weak scaling) and
one NCCL reduce sums the fp32 accumulation buffers on rank 0.

    python bench.py --gpus 1 --steps 5 --warmup 3            # this backend (CUDA)
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

PUBLISHED_MPATHS = 0.585  # BASELINE.md section 1 (README.md:13 of the reference: 1200x1200 @1000 spp in ~41 min, Apple M3 Pro)
B_RAY = 768  # algorithmic bytes per ray for C5: 160 state + 7 x 80 node + 48 triangle
SCENE = "cornell_dragon"


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


def gpu_batch_paths(npix):
    """Paths per batch of the render loop (api.cu: enqueue_render): whole images of samples, about 24 Mi paths."""
    mi = int(os.environ.get("RTGPU_BATCH_MPATHS", "24")) << 20
    return max(1, mi // npix) * npix


def trace_launches_of(paths, npix, max_depth, lanes, total_mem):
    """k_trace_fast launches of one render (api.cu: enqueue_render): every batch runs bounces < RTGPU_TAIL_BOUNCE, the last
    batch of each tail-merging group of a lane runs the others."""
    cap = min(paths, gpu_batch_paths(npix))
    n_batches = -(-paths // cap)
    lanes = max(1, min(lanes, n_batches))
    per_lane = -(-n_batches // lanes)
    group, bounce = max(1, int(os.environ.get("RTGPU_TAIL_GROUP", "3"))), int(os.environ.get("RTGPU_TAIL_BOUNCE", "4"))
    if bounce < 1 or bounce >= max_depth or per_lane < 2:
        group = 1
    group = min(group, per_lane)
    while group > 1 and lanes * cap * (304.0 + 216.0 * group) > int(os.environ.get("RTGPU_POOL_MEM_PCT", "60")) * 0.01 * total_mem:
        group -= 1
    if group == 1:
        return n_batches * max_depth
    total = 0
    for lane in range(lanes):
        nb = (n_batches - lane + lanes - 1) // lanes
        total += nb * bounce + -(-nb // group) * (max_depth - bounce)
    return total


def load_workload(api, width, replicas, sqrt_spt, max_depth=20):
    from tools.make_dragon_proxy import ensure
    ensure()
    spp = replicas * sqrt_spt * sqrt_spt
    scene, cam, prm = api.load([api.scene_path(SCENE), f"-w={width}", f"-t={replicas}", f"-s={spp}",
                                f"--max-depth={max_depth}", "--light-bias=0.25"])
    assert prm.replicas * prm.sqrt_spt ** 2 == spp
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
    scene, cam, prm = load_workload(api, args.width, 10, 2)
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
    sample = f"{sp.replicas} replicas x 1 stratum on every {row_step}-th row of 1200x1200 ({paths} paths/step, {cores} threads)"
    out = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": value / PUBLISHED_MPATHS, "dtype": "f64", "data": "synthetic (proxy dragon mesh)",
           "config": {"workload": f"scenes/{SCENE} (PROXY mesh, 1007616 tris) 1200x1200 max-depth 20 light-bias 0.25", "sample": sample},
           "mrays_per_s": rays / dt / 1e6,
           "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def run_ours(args):
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
    T, k = 10, args.sqrt_spt
    spp_rank = T * k * k
    if rank == 0:
        from tools.make_dragon_proxy import ensure
        ensure()
    if world > 1:
        dist.barrier()
    t_load = time.perf_counter()
    scene, cam, prm = load_workload(api, args.width, T * world, k)
    t_load = time.perf_counter() - t_load
    prm = sharding.shard_params(prm, rank, world)   # this rank's block of the T*k*k sample strata
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

    # ---- end to end through the host-buffer C ABI call (Camera::render drop-in) -----------------
    total_spp = spp_rank * world
    host = np.zeros((H, W, 4), dtype=np.float64)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    def e2e_step():
        if world == 1:
            gpu.render(cam, prm, out=host)          # rtgpu_render: kernels + D2H of the f64 Buffer, host add
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
    for _ in range(e2e_steps):
        e2e_step()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_value = paths_step_rank * world * e2e_steps / float(tt.item()) / 1e6

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
    rays_per_bounce = [int(sp1.rays_per_bounce[i] - sp0.rays_per_bounce[i]) for i in range(pprof.max_depth)]
    traversed_per_bounce = [int(sp1.rays_traversed_per_bounce[i] - sp0.rays_traversed_per_bounce[i]) for i in range(pprof.max_depth)]
    ms_trace, ms_shade, ms_gen = sp1.ms_trace - sp0.ms_trace, sp1.ms_shade - sp0.ms_shade, sp1.ms_raygen - sp0.ms_raygen
    trace_launches = trace_launches_of(paths_step_rank, W * H, pprof.max_depth, 1, torch.cuda.mem_get_info()[1])  # (profiled step: one lane)
    peak, peak_src = load_peaks()
    achieved = prof_rays * B_RAY / (ms_trace / 1e3) / 1e9 if ms_trace > 0 else None
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "trace_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_ray") * prof_rays / trace_launches
    except Exception:  # noqa: BLE001
        pass
    st = gpu.stats()
    out = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": value / PUBLISHED_MPATHS, "dtype": "f32",
        "vs_baseline_note": "BASELINE.md section 1: 0.585 Mpaths/s, derived from the reference README's ~41 min for this scene at 1000 spp on an "
                            "Apple M3 Pro (10 threads) with the real dragon mesh; here a proxy mesh of the same size, see `data`",
        "data": "synthetic (proxy dragon mesh; the reference's dragon_high.obj is missing upstream)",
        "config": {"workload": f"scenes/{SCENE} (PROXY mesh, {st.bvh_triangles} tris) {W}x{H} max-depth 20 light-bias 0.25, "
                               f"step = {spp_rank} spp per GPU in one render call ({T} replicas x {k}x{k} strata"
                               + (": the whole 1000-spp job)" if spp_rank == 1000 else ", a slice of the 1000-spp job)"),
                   "paths_per_step": paths_step_rank * world, "l2": "working set (path pools of ~23 GB per lane: streamed ray records, tail-merging carry arrays; scene ~200 MB) exceeds the 126 MB L2",
                   "parallelism": f"sample strata sharded over {world} GPU(s), one NCCL reduce of the fp32 accumulator per step"
                   if world > 1 else "1 GPU"},
        "mrays_per_s": rays_total / (ms / 1e3) / 1e6, "rays_per_path": rays_total / (paths_step_rank * world * args.steps),
        "kernel_ms_one_step_profiled": {"trace": ms_trace, "shade": ms_shade, "raygen": ms_gen,
                                        "note": "rank 0, one extra step on ONE lane with an event pair per launch; the timed "
                                                "region overlaps 4 batches on 4 streams and is shorter than this sum"},
        "rays_per_bounce_one_step": rays_per_bounce,
        "rays_traversed_per_bounce_one_step": traversed_per_bounce,
        "rays_resolved_by_broadcast_pretest": 1.0 - prof_rays / prof_rays_all if prof_rays_all > 0 else None,
        "roofline": {"bound": "hbm", "kernel": "k_trace_fast", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_ray": B_RAY, "launches": trace_launches,
                     "rays_per_launch": prof_rays / trace_launches,
                     "note": "per launch (mean over the %d k_trace_fast launches of one profiled step): achieved = rays the kernel traversed x 768 B / "
                             "CUDA-event duration on the launch stream, launches timed alone; traffic = ncu dram bytes per ray "
                             "(profiles/trace_traffic.json) x rays per launch, in bytes" % trace_launches},
        "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": C.sizeof(abi.CameraDesc) + C.sizeof(abi.RenderParams),
                "d2h_bytes_per_step": W * H * 4 * 8 if world == 1 else W * H * 4 * 8,
                "note": "rtgpu_render(host f64 Buffer): scene uploaded once beforehand like the reference's 'Ready' phase"},
        "gpu_launches": int(launches), "clocks": clocks,
        "build": {"bvh_nodes": st.bvh_nodes, "bvh_prims": st.bvh_prims, "ms_flatten_and_bvh": st.ms_build,
                  "ms_scene_dsl_and_obj_parse": t_load * 1e3, "ms_scene_create_incl_upload": t_create * 1e3,
                  "note": "the reference's 'Ready' phase, outside the timed region; multi-threaded host code"},
    }
    if world == 1 and not args.no_cpu:
        import oracle
        cores = os.cpu_count() or 1
        chk, sp, row_step = cpu_sample(oracle, scene, cam, prm, cores, 12.0)
        rows = len(range(0, H, row_step))
        t0 = time.perf_counter()
        _, r = chk.render_rows(cam, sp, 0, row_step, n_threads=cores)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": rows * W * sp.replicas / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                               "mrays_per_s": r / dt / 1e6,
                               "sample": f"{sp.replicas} replicas x 1 stratum on every {row_step}-th row of {W}x{H} ({rows * W * sp.replicas} paths, {dt:.1f} s, {cores} threads)"}
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1200)
    ap.add_argument("--sqrt-spt", type=int, default=10, help="strata per replica = sqrt_spt^2; spp per GPU per step = 10 * sqrt_spt^2 "
                    "(default 10: one step = the whole BASELINE job, 10 replicas x 10x10 strata = 1000 spp, in one render call per GPU)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
