"""N > 1 host logic on CPU: two gloo ranks each render their block of the sample strata (with
the oracle standing in for the device), all-reduce the buffers, and must reproduce the
single-process render (the reference sums its per-thread buffers the same way, camera.rs:243-255)."""
import os
import subprocess
import sys
import textwrap

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, {root!r})
    import numpy as np, torch, torch.distributed as dist
    from rust_raytracer_b200 import api, sharding
    import oracle
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    scene, cam, prm = api.load([api.scene_path("cornell"), "-w=48", "-s=64", "-t=4"])
    p = sharding.shard_params(prm, rank, world)
    assert (p.sample_end - p.sample_begin) * world == prm.replicas * prm.sqrt_spt ** 2
    buf, _ = oracle.OracleScene(scene).render(cam, p, n_threads=2)
    t = torch.from_numpy(buf)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.save({out!r}, t.numpy())
    dist.destroy_process_group()
""")


def test_two_rank_shards_sum_to_the_full_render(api, oracle, tmp_path):
    out = str(tmp_path / "sum.npy")
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, out=out))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", "29517", str(script)], env=env, timeout=600)
    got = np.load(out)
    scene, cam, prm = api.load([api.scene_path("cornell"), "-w=48", "-s=64", "-t=4"])
    full, _ = oracle.OracleScene(scene).render(cam, prm, n_threads=2)
    assert np.allclose(got, full, rtol=1e-12, atol=1e-14)


def test_shard_ranges_partition_the_samples():
    from rust_raytracer_b200 import sharding
    for total in (1, 7, 40, 1000):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(total, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
