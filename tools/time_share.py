#!/usr/bin/env python
"""tools/time_share.py — device time of ONE rank's share of the strong-scaling job on one GPU (GPU box only):
renders samples [0, S / N) of cornell_dragon 1200x1200 (S = 1000) for N in --ranks under the given env settings."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", default="1,2,4,8")
    ap.add_argument("--env", default="", help="semicolon-separated settings, each 'K=V,K=V'")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    import torch
    from rust_raytracer_b200 import api
    from tools.make_dragon_proxy import ensure
    ensure()
    scene, cam, prm = api.load([api.scene_path("cornell_dragon"), "-w=1200", "-t=10", "-s=1000", "--max-depth=20", "--light-bias=0.25"])
    gpu = api.GpuScene(scene, devices=[0])
    accum = torch.zeros((cam.height, cam.width, 4), dtype=torch.float32, device="cuda")
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    for setting in (args.env.split(";") if args.env else [""]):
        kv = dict(x.split("=") for x in setting.split(",") if x)
        for k, v in kv.items():
            os.environ[k] = v
        for n in [int(x) for x in args.ranks.split(",")]:
            p = type(prm).from_buffer_copy(prm)
            p.sample_begin, p.sample_end = 0, 1000 // n
            best = 1e30
            for _ in range(args.reps + 1):
                accum.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); gpu.render_device(cam, p, accum.data_ptr(), 0, stream.cuda_stream); e1.record(stream)
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            st = gpu.stats()
            print(json.dumps({"env": kv, "ranks": n, "spp": 1000 // n, "ms": round(best, 2), "ideal_ms_from_n1": None, "lanes": st.lanes,
                              "batch_paths": st.batch_paths, "tail_group": st.tail_group, "mpaths_s": round(1.44e6 * (1000 // n) / best / 1e3, 1)}), flush=True)
        for k in kv:
            os.environ.pop(k, None)


if __name__ == "__main__":
    main()
