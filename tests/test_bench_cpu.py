"""Host-side logic of bench.py (no GPU): workloads, synthetic camera, scene helpers."""
import importlib.util
import math
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_workloads_name_the_baseline_configs():
    b = load_bench()
    wl = b.WORKLOADS
    assert wl["smoke10m_1080p"]["photons"] == 10_000_000 and (wl["smoke10m_1080p"]["w"], wl["smoke10m_1080p"]["h"]) == (1920, 1080)
    assert wl["fog1m_1080p"]["grid"] == 0 and wl["fog1m_1080p"]["photons"] == 1_000_000      # configs[1]: homogeneous fog
    assert wl["smoke50m_4k"]["photons"] == 50_000_000 and (wl["smoke50m_4k"]["w"], wl["smoke50m_4k"]["h"]) == (3840, 2160)
    assert len(b.scene_lights(wl["fog1m_1080p"])) == 1 and len(b.scene_lights(wl["smoke10m_1080p"])) == 2
    # both arms print the same config keys (the driver compares them)
    assert set(b.base_config("smoke10m_1080p", wl["smoke10m_1080p"])) == set(b.base_config("fog1m_1080p", wl["fog1m_1080p"]))


def test_camera_matrices_describe_the_same_pinhole_as_camera_rays():
    """bre_camera of the end-to-end path (K0) vs the numpy rays of the device-resident path."""
    b = load_bench()
    w, h = 64, 36
    r2c, c2w = b.camera_matrices(w, h)
    _, d = b.camera_rays(w, h)
    ys, xs = np.meshgrid(np.arange(h) + 0.5, np.arange(w) + 0.5, indexing="ij")
    p = np.stack([xs.ravel(), ys.ravel(), np.zeros(w * h), np.ones(w * h)], axis=1) @ r2c.T.astype(np.float64)
    pc = p[:, :3] / p[:, 3:4]
    pc /= np.linalg.norm(pc, axis=1, keepdims=True)
    assert np.allclose(pc @ c2w[:3, :3].T.astype(np.float64), d, atol=1e-6)
    assert np.allclose(c2w[:3, 3], b.EYE)
    # per-rank sample offsets of the weak-scaling mode move the rays inside their pixels
    _, d2 = b.camera_rays(w, h, off=(0.25, 0.75))
    assert not np.array_equal(d, d2) and np.abs(d - d2).max() < 2.0 / h


def test_reference_arm_runs_the_unmodified_pbrt_binary():
    """bench.py --impl reference: one bounded run of oracle/_ref/pbrt (Integrator "photonbeam") per step."""
    import json
    import subprocess
    import sys
    import pytest
    b = load_bench()
    if not os.path.exists(b.PBRT_BIN):
        pytest.skip("oracle/_ref/pbrt has not been built (make -C oracle pbrt needs /root/reference)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--no-port", "--ref-photons", "3000", "--ref-crop", "0.06"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["cpu_baseline"]["kind"] == "reference" and line["value"] > 0
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    assert set(line["config"]) == set(b.base_config("smoke10m_1080p", b.WORKLOADS["smoke10m_1080p"]))
    assert line["cpu_baseline"]["pbrt_profile_s"]["camera_pass"] > 0 and "photonbeam" in line["cpu_baseline"]["sample"]


def test_camera_rays_are_unit_and_end_at_the_medium_box():
    b = load_bench()
    o, d = b.camera_rays(64, 36)
    assert o.shape == (64 * 36, 4) and d.shape == (64 * 36, 3) and o.dtype == np.float32
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-6)
    assert np.all(o[:, :3] == np.array([0.5, 0.5, -1.5], np.float32))
    hit = o[:, 3] > 0
    assert hit.mean() > 0.75                                 # the 40-degree camera looks into the box (edge rows pass above / below it)
    end = o[hit, :3] + d[hit] * o[hit, 3:4]
    assert np.all(end <= b.BOX_HI + 1e-4) and np.all(end >= b.BOX_LO - 1e-4)
    on_face = np.minimum(np.abs(end - b.BOX_LO), np.abs(end - b.BOX_HI)).min(axis=1)
    assert np.all(on_face < 1e-4)                            # tMax = exit point of the box
    # row windows tile the image
    o2, d2 = b.camera_rays(64, 36, 10, 20)
    assert np.array_equal(o2, o[10 * 64:20 * 64]) and np.array_equal(d2, d[10 * 64:20 * 64])


def test_smoke_density_is_deterministic_and_positive():
    b = load_bench()
    a, c = b.smoke_density(16), b.smoke_density(16)
    assert a.shape == (16, 16, 16) and a.dtype == np.float32 and np.array_equal(a, c)
    assert a.min() >= 0.02 - 1e-6 and a.max() > 0.5
    m = b.world_to_medium()
    lo = m @ np.append(b.BOX_LO, 1.0).astype(np.float32)
    hi = m @ np.append(b.BOX_HI, 1.0).astype(np.float32)
    assert np.allclose(lo[:3], 0.0, atol=1e-6) and np.allclose(hi[:3], 1.0, atol=1e-6)   # box -> unit cube
