"""Photon dump / load and film checkpoint files, through the C ABI on the GPU."""
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SA, SS = [0.05] * 3, [0.4] * 3


def read_photon_file(path):
    """Independent reader of the documented layout (include/bre.h)."""
    raw = open(path, "rb").read()
    assert raw[:8] == b"BREPHOT\0"
    version, zero, n, n_paths = struct.unpack_from("<IIQQ", raw, 8)
    assert version == 1 and zero == 0 and raw[32:64] == bytes(32)
    assert len(raw) == 64 + 48 * n
    a = np.frombuffer(raw, np.float32, offset=64).reshape(3, n, 4)
    return n_paths, a[0], a[1], a[2]


def test_photon_file_round_trip_and_layout(tmp_path, oracle, ctx, bre):
    n = 5000
    pos, pw, dr = oracle.synth_photons_uniform(n)
    radii = np.linspace(0.01, 0.03, n).astype(np.float32)
    ctx.set_medium_homogeneous(SA, SS, 0.3)
    ctx.upload_photons(pos, pw, dr, radii=radii)
    path = tmp_path / "photons.bin"
    ctx.save_photons(path, n_paths=1234)
    n_paths, a_pos, a_pow, a_dir = read_photon_file(path)
    assert n_paths == 1234
    assert np.array_equal(a_pos[:, :3], pos) and np.array_equal(a_pos[:, 3], radii)
    assert np.array_equal(a_pow[:, :3], pw) and np.array_equal(a_dir[:, :3], dr)
    # load into a second context: identical store, identical hits
    ro, rd = oracle.synth_rays_pinhole(32, 24)
    ctx.build()
    rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    c2 = bre.Bre(device=0)
    try:
        c2.set_medium_homogeneous(SA, SS, 0.3)
        assert c2.load_photons(path) == (n, 1234)
        for a, b in zip(c2.download_photons(), ctx.download_photons()):
            assert np.array_equal(a, b)
        c2.build()
        rgb2, offs2, idx2 = c2.query(ro, rd, float(n), want_hits=True)
        assert np.array_equal(offs, offs2) and np.array_equal(idx, idx2)
        assert np.allclose(rgb, rgb2, rtol=1e-5, atol=0)
        # append doubles the store
        assert c2.load_photons(path, append=True)[0] == n
        assert c2.photon_count() == 2 * n
    finally:
        c2.close()


def test_photon_file_errors(tmp_path, ctx, bre):
    ctx.set_medium_homogeneous(SA, SS, 0.0)
    ctx.upload_photons(np.zeros((10, 3), np.float32), np.ones((10, 3), np.float32), np.ones((10, 3), np.float32), radius=0.1)
    good = tmp_path / "p.bin"
    ctx.save_photons(good)
    raw = good.read_bytes()
    (tmp_path / "trunc.bin").write_bytes(raw[:-7])
    (tmp_path / "magic.bin").write_bytes(b"NOTPHOTO" + raw[8:])
    (tmp_path / "vers.bin").write_bytes(raw[:8] + struct.pack("<I", 9) + raw[12:])
    for name in ("trunc.bin", "magic.bin", "vers.bin", "missing.bin"):
        with pytest.raises(bre.BreError):
            ctx.load_photons(tmp_path / name)
        assert ctx.photon_count() == 10          # a failed load leaves the store alone
    ctx.clear_photons()
    ctx.save_photons(tmp_path / "empty.bin")
    assert ctx.load_photons(tmp_path / "empty.bin") == (0, 0)


def test_film_checkpoint_resume(tmp_path, oracle, ctx, bre):
    """Two iterations in one run == one iteration, checkpoint, second iteration in a new context."""
    n = 20000
    pos, pw, dr = oracle.synth_photons_uniform(n)
    ro, rd = oracle.synth_rays_pinhole(40, 30)
    ctx.set_medium_homogeneous(SA, SS, 0.3)

    def iteration(c, radius):
        c.upload_photons(pos, pw, dr, radius=radius)
        c.build()
        c.query_to_film(ro, rd, float(n))

    ctx.film_create(40 * 30)
    iteration(ctx, 0.02)
    ck = tmp_path / "film.ck"
    ctx.film_save(ck, 1, 0.015)
    iteration(ctx, 0.015)
    full = ctx.film_resolve(1, 1.0)

    c2 = bre.Bre(device=0)
    try:
        c2.set_medium_homogeneous(SA, SS, 0.3)
        c2.film_create(40 * 30)
        it, radius = c2.film_load(ck)
        assert it == 1 and radius == np.float32(0.015)
        iteration(c2, radius)
        resumed = c2.film_resolve(1, 1.0)
        assert np.allclose(resumed, full, rtol=1e-5, atol=1e-9 * full.max())
        # wrong resolution / truncated file: loud error, film untouched
        c2.film_create(10)
        with pytest.raises(bre.BreError):
            c2.film_load(ck)
        c2.film_create(40 * 30)
        (tmp_path / "short.ck").write_bytes(ck.read_bytes()[:-100])
        c2.film_accumulate(np.ones((40 * 30, 3), np.float32))
        with pytest.raises(bre.BreError):
            c2.film_load(tmp_path / "short.ck")
        assert np.allclose(c2.film_mean(0), 1.0)
    finally:
        c2.close()
