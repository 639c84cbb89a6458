"""GPU parity for the surface rows of SURVEY.md 8f-1 / 8f-2, through the C ABI.

Golden = tests/golden/ref_surfaces.npz: photon paths walked by REFERENCE classes through a fog box
with matte triangle walls, one reverse-orientation wall and an interface-only medium boundary, plus
camera rays of the reference's PerspectiveCamera with first-hit tMax (oracle/ref_harness.cpp
`surfaces`).  Camera rays are pure float arithmetic -> bit-exact.  Photon paths use the GPU's own
logf/expf/sinf/cosf -> float tolerance path by path, like the volume-only tracing test."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLD, "ref_surfaces.npz"))


def setup_scene(ctx, g):
    ctx.set_medium_homogeneous(g["sigma_a"], g["sigma_s"], float(g["g"][0]))
    ctx.set_triangles(g["mesh_P"], g["mesh_idx"], g["mesh_Kd"], g["mesh_attr"].astype(np.uint8))
    ctx.set_lights([dict(kind=0, pos=g["light_pos"], I=g["light_I"])])


def matching_paths(pos, pw, dr, rp, ra, rw, rid):
    """Photons are stored in path order on both sides; count reference paths reproduced photon by
    photon within tolerance, resynchronising on the next path after a mismatch."""
    n = pos.shape[0]
    same = i = 0
    starts = np.flatnonzero(np.r_[True, rid[1:] != rid[:-1]])
    ends = np.r_[starts[1:], len(rid)]
    for s, e in zip(starts, ends):
        k = e - s
        if i + k <= n and np.allclose(pos[i:i + k], rp[s:e], atol=2e-4) and \
                np.allclose(pw[i:i + k], ra[s:e], rtol=3e-3) and np.allclose(dr[i:i + k], rw[s:e], atol=2e-4):
            same += 1
            i += k
            continue
        if e >= len(rid):
            break
        d = np.abs(pos[i:i + 96] - rp[e]).max(axis=1)
        hit = np.flatnonzero(d < 2e-4)
        if hit.size == 0:
            break
        i += int(hit[0])
    return same, len(starts)


def test_camera_rays_bit_exact_vs_reference_camera(ctx, g):
    setup_scene(ctx, g)
    x0, y0, w, h = [int(v) for v in g["window"]]
    ctx.set_camera(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h)
    ro, rd = ctx.camera_rays(float(g["pixel_offset"][0]), float(g["pixel_offset"][1]))
    assert np.array_equal(bits(ro), bits(g["ray_o_tmax"].reshape(-1, 4)))
    assert np.array_equal(bits(rd), bits(g["ray_d"].reshape(-1, 3)))


def test_camera_rays_without_mesh_and_outside_medium(ctx, g, oracle):
    ctx.set_medium_homogeneous(g["sigma_a"], g["sigma_s"], float(g["g"][0]))
    x0, y0, w, h = [int(v) for v in g["window"]]
    ctx.set_camera(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h)
    ro, rd = ctx.camera_rays(0.5, 0.5)
    oracle.set_mesh(g["mesh_P"], g["mesh_idx"], g["mesh_Kd"])
    oo, od = oracle.camera_rays(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h, 0.5, 0.5, use_mesh=False)
    assert np.array_equal(bits(ro), bits(oo)) and np.array_equal(bits(rd), bits(od))
    assert (ro[:, 3] == np.finfo(np.float32).max).all()   # no surface: tMax = min(inf, MaxFloat)
    ctx.set_camera(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h, in_medium=False)
    ro, _ = ctx.camera_rays(0.5, 0.5)
    assert (ro[:, 3] == 0).all()             # camera outside any medium: nothing to gather


def test_device_surface_photon_tracing_vs_reference_classes(ctx, g):
    """K1 with triangles: Lambertian bounces, roulette, interface pass-through, medium bookkeeping."""
    setup_scene(ctx, g)
    first, n_paths, depth = [int(x) for x in g["range"]]
    n = ctx.trace_photons(first, n_paths, depth, 0.1)
    pos, pw, dr, _ = ctx.download_photons()
    rp, ra, rw = g["photon_p"].reshape(-1, 3), g["photon_alpha"].reshape(-1, 3), g["photon_wi"].reshape(-1, 3)
    assert abs(n - rp.shape[0]) <= 0.01 * rp.shape[0]
    same, total = matching_paths(pos, pw, dr, rp, ra, rw, g["path_id"])
    assert same >= 0.95 * total, f"{same}/{total} paths match"
    assert pos[:, 0].max() < 1.5 + 1e-4      # nothing is deposited beyond the interface (vacuum side)


def test_device_surface_tracing_matches_oracle_statistics(ctx, g, oracle):
    """Larger run against the oracle (which is bit-exact to the reference on the golden paths):
    photon count, mean power and spatial moments agree statistically."""
    setup_scene(ctx, g)
    oracle.set_medium_homogeneous(g["sigma_a"], g["sigma_s"], float(g["g"][0]))
    oracle.set_mesh(g["mesh_P"], g["mesh_idx"], g["mesh_Kd"], g["mesh_attr"].astype(np.uint8))
    n_paths, depth = 60000, 8
    n = ctx.trace_photons(5000, n_paths, depth, 0.1)
    pos, pw, dr, _ = ctx.download_photons()
    op, ow, od, _ = oracle.trace_mesh([dict(kind=0, pos=g["light_pos"], I=g["light_I"])], 5000, n_paths, depth)
    assert abs(n - len(op)) <= 0.003 * len(op)
    assert np.allclose(pos.mean(0), op.mean(0), atol=0.02)
    assert np.allclose(pos.std(0), op.std(0), rtol=0.01)
    assert np.allclose(pw.sum(0), ow.sum(0), rtol=0.005)


def test_camera_query_to_film_equals_host_ray_path(ctx, g, bre):
    """bre_query_camera_to_film (rays generated on the device) == bre_query_to_film fed with the
    same rays from the host: identical film, bit for bit is not required (atomics), 1e-6 is."""
    setup_scene(ctx, g)
    x0, y0, w, h = [int(v) for v in g["window"]]
    n_paths = 20000
    ctx.trace_photons(0, n_paths, 6, 0.15)
    ctx.build()
    ctx.set_camera(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h)
    ctx.film_create(w * h)
    ctx.query_camera_to_film(float(n_paths), 0.5, 0.5)
    a = ctx.film_mean(0)
    ro, rd = ctx.camera_rays(0.5, 0.5)
    ctx.film_clear()
    ctx.query_to_film(ro, rd, float(n_paths))
    b = ctx.film_mean(0)
    assert a.max() > 0
    assert np.allclose(a, b, rtol=1e-5, atol=1e-9 * a.max())


def test_triangle_api_errors(ctx, bre):
    with pytest.raises(bre.BreError):
        ctx.camera_rays()                    # no camera yet
    ctx.set_medium_homogeneous([0.1] * 3, [0.4] * 3, 0.0)
    ctx.set_lights([dict(kind=0, pos=(0, 0, 0), I=(1, 1, 1))])
    with pytest.raises(bre.BreError):
        ctx.trace_photons(0, 10, 3, 0.1)     # neither triangles nor bounds
    P = np.array([[0, 0, 1], [1, 0, 1], [0, 1, 1]], np.float32)
    ctx.set_triangles(P, [0, 1, 2], [0.5, 0.5, 0.5])
    assert ctx.trace_photons(0, 100, 3, 0.1) >= 0   # paths escape past a single triangle: no hang
    ctx.set_triangles(P, np.zeros(0, np.uint32), np.zeros(0, np.float32))   # n_tri = 0 removes the mesh
    with pytest.raises(bre.BreError):
        ctx.trace_photons(0, 10, 3, 0.1)


@pytest.mark.parametrize("sample", [0, 3, 1000])
def test_camera_rays_halton_samples_bit_exact(ctx, g, oracle, sample):
    """K0 with the reference's per-pixel Halton camera samples (bre_camera_rays_halton): bit-exact against the oracle's
    camera rays through the oracle's Halton sample positions -- which tests/test_oracle_golden.py pins, bit for bit,
    to the reference's HaltonSampler (tests/golden/ref_halton.npz) and PerspectiveCamera (ref_surfaces.npz)."""
    setup_scene(ctx, g)
    x0, y0, w, h = [int(v) for v in g["window"]]
    ctx.set_camera(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h)
    ro, rd = ctx.camera_rays_halton(sample)
    offs = oracle.halton_camera_samples(x0, y0, x0 + w, y0 + h, sample)
    oracle.set_mesh(g["mesh_P"], g["mesh_idx"], g["mesh_Kd"], g["mesh_attr"].astype(np.uint8))
    oo, od = oracle.camera_rays_px(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h, offs)
    assert np.array_equal(bits(ro), bits(oo))
    assert np.array_equal(bits(rd), bits(od))
    # not the pixel centres, and not one offset for every pixel
    c_ro, c_rd = ctx.camera_rays(0.5, 0.5)
    assert not np.array_equal(bits(rd), bits(c_rd))
    assert len(np.unique(offs[:, 0])) > min(w, 16)


def test_camera_halton_query_to_film_staged_equals_unstaged(ctx, g):
    """bre_stage_camera_halton + bre_query_camera_halton_to_film (K0 / K6 under the photon pass) == the unstaged call
    == bre_query_to_film fed with the rays of bre_camera_rays_halton; another sample number is not picked up."""
    setup_scene(ctx, g)
    x0, y0, w, h = [int(v) for v in g["window"]]
    n_paths = 20000
    ctx.set_camera(g["raster_to_camera"], g["camera_to_world"], x0, y0, w, h)
    ctx.film_create(w * h)
    ctx.stage_camera_halton(5)
    ctx.trace_photons(0, n_paths, 6, 0.15)
    ctx.build()
    ctx.query_camera_halton_to_film(float(n_paths), 5)
    a = ctx.film_mean(0)
    ctx.film_clear()
    ctx.query_camera_halton_to_film(float(n_paths), 5)          # nothing staged now
    b = ctx.film_mean(0)
    ro, rd = ctx.camera_rays_halton(5)
    ctx.film_clear()
    ctx.query_to_film(ro, rd, float(n_paths))
    c = ctx.film_mean(0)
    ctx.film_clear()
    ctx.stage_camera_halton(5)
    ctx.query_camera_halton_to_film(float(n_paths), 6)          # staged sample 5, asked for 6: regenerated
    d = ctx.film_mean(0)
    assert a.max() > 0
    assert np.allclose(a, b, rtol=1e-5, atol=1e-9 * a.max())
    assert np.allclose(a, c, rtol=1e-5, atol=1e-9 * a.max())
    assert not np.allclose(a, d, rtol=1e-5, atol=1e-9 * a.max())
