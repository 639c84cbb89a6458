"""GPU parity tests: libbre.so (through the C ABI) against the CPU oracle on the same inputs.

Bar: photon hit-index sets per ray bit-exact; canonical counters V/T/H bit-exact; the canonical
LBVH (sort order, Morton codes, topology, node boxes) bit-exact; radiance within 1e-4 mean /
1e-3 max relative (BASELINE.json north_star).
"""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SA, SS = [0.05, 0.05, 0.05], [0.4, 0.4, 0.4]
MEAN_TOL, MAX_TOL = 1e-4, 1e-3   # relative radiance tolerance stated by north_star


def radius_for(n, h):
    return math.sqrt(h / (math.pi * n))


def assert_radiance(rgb, ref):
    err = np.abs(rgb.astype(np.float64) - ref)
    assert err.mean() <= MEAN_TOL * max(np.abs(ref).mean(), 1e-30)
    assert err.max() <= MAX_TOL * max(np.abs(ref).max(), 1e-30)


def setup_uniform(oracle, ctx, n, h, g=0.3, radii=None):
    pos, pw, dr = oracle.synth_photons_uniform(n)
    r = radius_for(n, h) if radii is None else radii
    oracle.set_medium_homogeneous(SA, SS, g)
    oracle.set_photons(pos, pw, dr, r)
    oracle.build()
    ctx.set_medium_homogeneous(SA, SS, g)
    if radii is None:
        ctx.upload_photons(pos, pw, dr, radius=r)
    else:
        ctx.upload_photons(pos, pw, dr, radii=radii)
    ctx.build()
    return pos, pw, dr


@pytest.mark.parametrize("n", [2, 3, 1000, 50000])
def test_canonical_tree_bit_exact(oracle, ctx, n):
    setup_uniform(oracle, ctx, n, 16)
    o_order, o_codes, o_lo, o_hi, o_ch = oracle.tree_arrays()
    order, codes, lo, hi, ch = ctx.download_tree()
    assert np.array_equal(order, o_order)
    assert np.array_equal(codes, o_codes)
    assert np.array_equal(ch, o_ch)
    assert np.array_equal(lo.view(np.uint32), o_lo.view(np.uint32))
    assert np.array_equal(hi.view(np.uint32), o_hi.view(np.uint32))


@pytest.mark.parametrize("rays", ["pinhole", "incoherent"])
@pytest.mark.parametrize("n,h", [(20000, 8), (100000, 64)])
def test_hit_sets_and_radiance(oracle, ctx, n, h, rays):
    setup_uniform(oracle, ctx, n, h)
    ro, rd = oracle.synth_rays_pinhole(64, 48) if rays == "pinhole" else oracle.synth_rays_incoherent(3000)
    ref_rgb, ref_offs, ref_idx, ref_cnt = oracle.query(ro, rd, float(n), want_hits=True)
    rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    assert ref_offs[-1] > 0
    assert np.array_equal(offs, ref_offs)
    assert np.array_equal(idx, ref_idx)
    assert_radiance(rgb, ref_rgb)
    cnt = ctx.count(ro, rd)
    for k in ("nodes_visited", "disk_tests", "hits"):
        assert cnt[k] == ref_cnt[k], k
    q = ro.shape[0]
    assert cnt["bytes_algorithmic"] == 48 * q + 32 * cnt["nodes_visited"] + 16 * cnt["disk_tests"] + 32 * cnt["hits"]


def test_brute_force_agrees(oracle, ctx):
    """The tree never culls a true hit: GPU hit sets == brute force over all photons."""
    n = 30000
    setup_uniform(oracle, ctx, n, 32)
    ro, rd = oracle.synth_rays_incoherent(500)
    _, b_offs, b_idx, _ = oracle.query(ro, rd, float(n), want_hits=True, brute=True)
    _, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    assert np.array_equal(offs, b_offs) and np.array_equal(idx, b_idx)


@pytest.mark.parametrize("leaf", [1, 4, 16])
def test_leaf_sizes(oracle, bre, leaf):
    c = bre.Bre(device=0, max_leaf_photons=leaf)
    try:
        n = 20000
        setup_uniform(oracle, c, n, 16)
        ro, rd = oracle.synth_rays_pinhole(40, 40)
        ref_rgb, ref_offs, ref_idx, _ = oracle.query(ro, rd, float(n), want_hits=True)
        rgb, offs, idx = c.query(ro, rd, float(n), want_hits=True)
        assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
        assert_radiance(rgb, ref_rgb)
    finally:
        c.close()


def test_ragged_radii_and_collisions(oracle, ctx):
    """Per-photon radii, many photons at identical positions (Morton-code collisions)."""
    n = 8000
    rng = np.random.default_rng(7)
    pos, pw, dr = oracle.synth_photons_uniform(n)
    pos[1000:3000] = pos[1000]            # 2000 coincident photons
    pos[3000:3500, 0] = pos[3000, 0]      # a plane of equal x
    radii = (radius_for(n, 16) * rng.uniform(0.25, 3.0, n)).astype(np.float32)
    oracle.set_medium_homogeneous(SA, SS, 0.0)
    oracle.set_photons(pos, pw, dr, radii)
    oracle.build()
    ctx.set_medium_homogeneous(SA, SS, 0.0)
    ctx.upload_photons(pos, pw, dr, radii=radii)
    ctx.build()
    o_tree = oracle.tree_arrays()
    g_tree = ctx.download_tree()
    for a, b in zip(g_tree, o_tree):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    ro, rd = oracle.synth_rays_pinhole(32, 32)
    ref_rgb, ref_offs, ref_idx, ref_cnt = oracle.query(ro, rd, float(n), want_hits=True)
    rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
    assert_radiance(rgb, ref_rgb)
    cnt = ctx.count(ro, rd)
    assert cnt["nodes_visited"] == ref_cnt["nodes_visited"]


def test_empty_and_tiny(oracle, ctx):
    ctx.set_medium_homogeneous(SA, SS, 0.0)
    ro, rd = oracle.synth_rays_pinhole(8, 8)
    # no photons at all
    ctx.clear_photons()
    ctx.build()
    rgb, offs, idx = ctx.query(ro, rd, 1.0, want_hits=True)
    assert not rgb.any() and offs[-1] == 0 and idx.size == 0
    # no rays
    rgb = ctx.query(np.zeros((0, 4), np.float32), np.zeros((0, 3), np.float32), 1.0)
    assert rgb.shape == (0, 3)
    # one photon on the optical axis
    pos = np.array([[0.5, 0.5, 0.5]], np.float32)
    pw = np.ones((1, 3), np.float32); dr = np.array([[0, 0, 1]], np.float32)
    oracle.set_medium_homogeneous(SA, SS, 0.0)
    oracle.set_photons(pos, pw, dr, 0.2)
    oracle.build()
    ctx.upload_photons(pos, pw, dr, radius=0.2)
    ctx.build()
    ref_rgb, ref_offs, ref_idx, _ = oracle.query(ro, rd, 1.0, want_hits=True)
    rgb, offs, idx = ctx.query(ro, rd, 1.0, want_hits=True)
    assert ref_offs[-1] > 0
    assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
    assert_radiance(rgb, ref_rgb)
    # rays with tMax = 0 never hit
    ro0 = ro.copy(); ro0[:, 3] = 0
    rgb, offs, idx = ctx.query(ro0, rd, 1.0, want_hits=True)
    assert offs[-1] == 0 and not rgb.any()


def test_query_before_build_fails_loudly(oracle, ctx, bre):
    ctx.set_medium_homogeneous(SA, SS, 0.0)
    pos, pw, dr = oracle.synth_photons_uniform(10)
    ctx.upload_photons(pos, pw, dr, radius=0.1)
    ro, rd = oracle.synth_rays_pinhole(2, 2)
    with pytest.raises(bre.BreError):
        ctx.query(ro, rd, 1.0)


def smoke_grid(nz=16, ny=16, nx=16):
    z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    c = np.array([nx, ny, nz]) / 2.0
    d = np.exp(-((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) / (0.08 * nx * nx)) + 0.05
    return d.astype(np.float32)


def test_grid_medium_radiance(oracle, ctx):
    n = 20000
    pos, pw, dr = oracle.synth_photons_uniform(n)
    r = radius_for(n, 32)
    dens = smoke_grid()
    w2m = np.eye(4, dtype=np.float32)
    sa, ss = [0.5] * 3, [3.5] * 3
    oracle.set_medium_grid(sa, ss, 0.3, dens, w2m)
    oracle.set_photons(pos, pw, dr, r)
    oracle.build()
    ctx.set_medium_grid(sa, ss, 0.3, dens, w2m)
    ctx.upload_photons(pos, pw, dr, radius=r)
    ctx.build()
    ro, rd = oracle.synth_rays_pinhole(48, 48)
    ref_rgb, ref_offs, ref_idx, _ = oracle.query(ro, rd, float(n), want_hits=True)
    rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
    assert_radiance(rgb, ref_rgb)


def test_film_parity(oracle, ctx):
    rng = np.random.default_rng(3)
    npx = 64 * 32
    ctx.film_create(npx)
    acc = np.zeros((npx, 3), np.float32)
    for it in range(3):
        rgb = rng.uniform(0, 2, (npx, 3)).astype(np.float32)
        ctx.film_accumulate(rgb)
        acc += rgb
    img = ctx.film_resolve(2, 0.5)
    ref = oracle.film_resolve(acc, 2, 0.5)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
    # scattered accumulation through pixel indices (several rays per pixel)
    ctx.film_clear()
    pix = rng.integers(0, npx, 5000).astype(np.uint32)
    rgb = rng.uniform(0, 1, (5000, 3)).astype(np.float32)
    ctx.film_accumulate(rgb, pix)
    ref_acc = np.zeros((npx, 3), np.float64)
    np.add.at(ref_acc, pix, rgb)
    img = ctx.film_resolve(0, 1.0)
    assert np.allclose(img, oracle.film_resolve(ref_acc.astype(np.float32), 0, 1.0), rtol=1e-5, atol=1e-6)


def test_trace_photons_statistical(oracle, ctx):
    """K1: device logf/sinf/cosf differ from glibc in the last ulp, so parity is statistical
    plus 'almost all records identical to 1e-5' (same PCG32 streams, same deposit rule)."""
    lights = [dict(kind=0, pos=(0.5, 0.9, 0.5), I=(5, 5, 5))]
    blo, bhi = (0, 0, 0), (1, 1, 1)
    sa, ss, g = [0.1] * 3, [1.9] * 3, 0.3
    oracle.set_medium_homogeneous(sa, ss, g)
    ctx.set_medium_homogeneous(sa, ss, g)
    ctx.set_lights(lights)
    ctx.set_bounds(blo, bhi)
    npaths = 20000
    n = ctx.trace_photons(0, npaths, 5, 0.05)
    pos, pw, dr, rad = ctx.download_photons()
    o_pos, o_pw, o_dr, _ = oracle.trace(lights, blo, bhi, 0, npaths, 5)
    assert n == pos.shape[0]
    assert abs(n - o_pos.shape[0]) <= max(3, int(1e-3 * n))
    assert np.all(rad == np.float32(0.05))
    if n == o_pos.shape[0]:
        close = np.all(np.abs(pos - o_pos) < 1e-4, axis=1)
        assert close.mean() > 0.99
    assert np.allclose(pos.mean(0), o_pos.mean(0), atol=5e-3)
    assert np.allclose(pw.sum(0), o_pw.sum(0), rtol=5e-3)
    assert np.all((pos >= 0) & (pos <= 1))
    assert np.allclose(np.linalg.norm(dr, axis=1), 1, atol=1e-4)
    # determinism + path-index order: tracing in two halves appends the same store
    ctx.clear_photons()
    ctx.trace_photons(0, npaths // 2, 5, 0.05)
    ctx.trace_photons(npaths // 2, npaths - npaths // 2, 5, 0.05)
    pos2, pw2, dr2, _ = ctx.download_photons()
    assert np.array_equal(pos2, pos) and np.array_equal(pw2, pw) and np.array_equal(dr2, dr)


@pytest.mark.parametrize("cap", [0, 1000])
def test_pair_queue_overflow_path(oracle, bre, monkeypatch, cap):
    """K5a hands (ray, leaf) pairs to K5b through a queue; with a queue that is far too small
    (cap=1000 pairs) the overflow pairs are shaded in place -- results must not change."""
    if cap:
        monkeypatch.setenv("BRE_PAIR_QUEUE_CAP", str(cap))
    c = bre.Bre(device=0)
    try:
        n = 40000
        setup_uniform(oracle, c, n, 48)
        ro, rd = oracle.synth_rays_pinhole(50, 37)      # ragged: not a multiple of the packet size
        ref_rgb, ref_offs, ref_idx, _ = oracle.query(ro, rd, float(n), want_hits=True)
        rgb, offs, idx = c.query(ro, rd, float(n), want_hits=True)
        assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
        assert_radiance(rgb, ref_rgb)
        cnt = c.count(ro, rd)
        assert (cnt["pair_queue_overflow"] > 0) == bool(cap)
    finally:
        c.close()


def narrow_pinhole(w, h, pitch=7e-4):
    """Rays of a pinhole at (0.5, 0.5, -1.5) through a w x h window of `pitch` rad pixels."""
    x, y = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    d = np.stack([(x - 0.5 * (w - 1)) * pitch, (y - 0.5 * (h - 1)) * pitch * 40, np.ones_like(x)], -1).reshape(-1, 3)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    o = np.empty((w * h, 4), np.float32)
    o[:, :3] = (0.5, 0.5, -1.5)
    o[:, 3] = 3.0
    return o, d


@pytest.mark.parametrize("cap", [0, 3000])
@pytest.mark.parametrize("spread", ["default", "0.5", "0"])
@pytest.mark.parametrize("leaf", [4, 8, 16])
def test_coherent_packets_walked_as_one_fat_ray(oracle, bre, monkeypatch, spread, leaf, cap):
    """Neighbouring camera rays are walked once per packet of 8 (interval slab test) and resolved
    per ray by the exact disk predicate: hit sets must not depend on whether a packet took the fat
    walk (default threshold, a huge threshold that makes every same-octant packet fat, or off),
    also through the in-place overflow path."""
    if spread != "default":
        monkeypatch.setenv("BRE_FAT_SPREAD", spread)
    if cap:
        monkeypatch.setenv("BRE_PAIR_QUEUE_CAP", str(cap))
    c = bre.Bre(device=0, max_leaf_photons=leaf)
    try:
        n = 60000
        setup_uniform(oracle, c, n, 32)
        ro, rd = narrow_pinhole(1029, 3)                # 7e-4 rad between neighbours, ragged tail
        ro[100:108, 3] = 0.0                            # a dead packet
        ro[200:204, 3] = 0.0                            # and a half-dead one
        ref_rgb, ref_offs, ref_idx, _ = oracle.query(ro, rd, float(n), want_hits=True)
        rgb, offs, idx = c.query(ro, rd, float(n), want_hits=True)
        assert ref_offs[-1] > 0
        assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
        assert_radiance(rgb, ref_rgb)
        cnt = c.count(ro, rd)
        assert (cnt["pair_queue_overflow"] > 0) == bool(cap)
    finally:
        c.close()


def test_heavy_ray_skew(oracle, ctx):
    """A ray through a dense photon cluster (light source in view): thousands of hits on one ray."""
    rng = np.random.default_rng(11)
    n = 60000
    pos, pw, dr = oracle.synth_photons_uniform(n)
    pos[:30000] = (np.array([0.5, 0.5, 0.5]) + rng.normal(0, 0.004, (30000, 3))).astype(np.float32)
    r = 0.01
    oracle.set_medium_homogeneous(SA, SS, 0.3)
    oracle.set_photons(pos, pw, dr, r)
    oracle.build()
    ctx.set_medium_homogeneous(SA, SS, 0.3)
    ctx.upload_photons(pos, pw, dr, radius=r)
    ctx.build()
    ro, rd = oracle.synth_rays_pinhole(33, 33)
    ref_rgb, ref_offs, ref_idx, _ = oracle.query(ro, rd, float(n), want_hits=True)
    assert np.diff(ref_offs.astype(np.int64)).max() > 5000
    rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
    assert_radiance(rgb, ref_rgb)


@pytest.mark.parametrize("n", [1, 5, 30000])
def test_progressive_radius_refit_only(oracle, ctx, n):
    """after bre_set_radius (r <- r*(i+alpha)/(i+1), integrators/photonbeam.cpp:354-356,562)
    bre_build keeps order and topology and only refits.  The refitted tree must be bit-identical to
    a full build at the new radius (= the oracle's), hit sets and radiance must follow the new radius."""
    pos, pw, dr = setup_uniform(oracle, ctx, n, 24)
    assert ctx.last_build_kind() == 1
    r = radius_for(n, 24)
    ro, rd = oracle.synth_rays_pinhole(40, 30)
    for it in range(3):
        r = r * (it + 0.7) / (it + 1)
        ctx.set_radius(r)
        with pytest.raises(Exception):
            ctx.query(ro, rd, float(n))          # stale tree: loud failure, not stale hits
        ctx.build()
        assert ctx.last_build_kind() == 2
        oracle.set_photons(pos, pw, dr, r)
        oracle.build()
        for a, b in zip(ctx.download_tree(), oracle.tree_arrays()):
            assert np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
        ref_rgb, ref_offs, ref_idx, ref_cnt = oracle.query(ro, rd, float(n), want_hits=True)
        rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
        assert np.array_equal(offs, ref_offs) and np.array_equal(idx, ref_idx)
        assert_radiance(rgb, ref_rgb)
        cnt = ctx.count(ro, rd)
        assert cnt["nodes_visited"] == ref_cnt["nodes_visited"] and cnt["disk_tests"] == ref_cnt["disk_tests"]
    # changing the photons drops the topology: full build again
    ctx.upload_photons(pos, pw, dr, radius=r)
    ctx.build()
    assert ctx.last_build_kind() == 1


def test_staged_rays_equal_direct_upload(oracle, ctx, bre):
    """bre_stage_rays (async upload on the copy stream, issued BEFORE the photons exist) followed by
    bre_query_staged_to_film gives the film of bre_query_to_film; a second query without staging fails."""
    n = 30000
    pos, pw, dr = oracle.synth_photons_uniform(n)
    ro, rd = oracle.synth_rays_pinhole(48, 36)
    ctx.set_medium_homogeneous(SA, SS, 0.3)
    ctx.film_create(48 * 36)
    ctx.stage_rays(ro, rd)                       # rays first ...
    ctx.upload_photons(pos, pw, dr, radius=radius_for(n, 32))
    ctx.build()                                  # ... photon structure afterwards
    ctx.query_staged_to_film(float(n))
    a = ctx.film_mean(0)
    ctx.film_clear()
    ctx.query_to_film(ro, rd, float(n))
    b = ctx.film_mean(0)
    assert a.max() > 0 and np.allclose(a, b, rtol=1e-5, atol=1e-9 * a.max())
    with pytest.raises(bre.BreError):
        ctx.query_staged_to_film(float(n))
