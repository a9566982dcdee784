"""Committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py from the
oracle): the oracle must keep reproducing them bit for bit; the GPU tests reuse the same rays."""
import os

import numpy as np
import pytest

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["cornell", "light_test", "perlin"])
def test_oracle_reproduces_golden_first_hits(api, oracle, name):
    g = np.load(os.path.join(HERE, f"{name}_first_hit.npz"))
    scene, cam, prm = api.load([api.scene_path(name)])
    o = oracle.OracleScene(scene)
    hits = o.trace_primary(g["rays"])
    for f in ("object", "instance", "primitive", "front_face"):
        assert np.array_equal(hits[f], g["hits"][f]), f
    assert np.array_equal(hits["t"], g["hits"]["t"])
    assert np.allclose(hits["normal"], g["hits"]["normal"], rtol=1e-12, atol=1e-15)
    b = o.bounce(g["rays"], g["xi"], 0.25)
    assert np.array_equal(b["kind"], g["bounce"]["kind"])
    assert np.allclose(b["next_dir"], g["bounce"]["next_dir"], rtol=1e-12, atol=1e-15)
    fin = np.isfinite(g["bounce"]["weight"]).all(axis=1)
    assert np.allclose(b["weight"][fin], g["bounce"]["weight"][fin], rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cornell", "light_test", "perlin"])
def test_gpu_reproduces_golden_first_hits(api, name):
    g = np.load(os.path.join(HERE, f"{name}_first_hit.npz"))
    scene, cam, prm = api.load([api.scene_path(name)])
    gpu = api.GpuScene(scene)
    hits = gpu.trace_primary(g["rays"])
    for f in ("object", "instance", "primitive", "front_face"):
        assert np.array_equal(hits[f], g["hits"][f]), f
    hit = g["hits"]["object"] >= 0
    assert np.array_equal(hits["t"][hit], g["hits"]["t"][hit])
