"""Drop-in check on the GPU: the pbrt-side adapter (host/bre_integrator.cpp), compiled against the
unmodified reference headers and linked with the reference's Scene/Camera/Film code and libbre.so
(oracle/_ref/bre_pbrt_render, built where /root/reference is mounted and shipped as a binary),
renders the fog box; the image the reference's Film wrote must equal the CPU oracle's estimate on
the very photons and rays the run used."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "bre_pbrt_render")
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = [int(x) for x in f.readline().split()]
        assert float(f.readline()) < 0          # little endian
        a = np.frombuffer(f.read(w * h * 12), dtype="<f4").reshape(h, w, 3)
    return a[::-1].reshape(-1, 3).copy()        # PFM scanlines run bottom to top


needs_bin = pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/bre_pbrt_render not built (needs /root/reference)")


def fog_box_mesh():
    """The 12 matte triangles bre_pbrt_render builds, in the same order."""
    L, H = (-4.0, -1.0, -4.0), (4.0, 2.0, 4.0)
    quads = [
        [(L[0], L[1], L[2]), (H[0], L[1], L[2]), (H[0], L[1], H[2]), (L[0], L[1], H[2])],
        [(L[0], H[1], L[2]), (L[0], H[1], H[2]), (H[0], H[1], H[2]), (H[0], H[1], L[2])],
        [(L[0], L[1], H[2]), (H[0], L[1], H[2]), (H[0], H[1], H[2]), (L[0], H[1], H[2])],
        [(L[0], L[1], L[2]), (L[0], H[1], L[2]), (H[0], H[1], L[2]), (H[0], L[1], L[2])],
        [(L[0], L[1], L[2]), (L[0], L[1], H[2]), (L[0], H[1], H[2]), (L[0], H[1], L[2])],
        [(H[0], L[1], L[2]), (H[0], H[1], L[2]), (H[0], H[1], H[2]), (H[0], L[1], H[2])]]
    P = np.asarray(quads, np.float32).reshape(-1, 3)
    idx = np.concatenate([[4 * q + k for k in (0, 1, 2, 0, 2, 3)] for q in range(6)]).astype(np.uint32)
    Kd = np.full((12, 3), 0.6, np.float32)
    return P, idx, Kd


@needs_bin
@pytest.mark.parametrize("tracer,camrays", [("host", "host"), ("host", "device"), ("device", "device")])
def test_pbrt_adapter_renders_fog_box(tmp_path, oracle, bre, tracer, camrays):
    from make_golden import read_bref
    pfm, dump = str(tmp_path / "out.pfm"), str(tmp_path / "dump.bin")
    res, n_paths, radius = 32, 6000, 0.2
    subprocess.run([BIN, pfm, dump, tracer, str(res), str(n_paths), str(radius), "1", camrays], check=True, timeout=300)
    img = read_pfm(pfm)
    d = read_bref(dump)
    m = d["medium"]
    assert int(d["used_device_camera"][0]) == (camrays == "device")
    ro, rd = d["ray_o_tmax"].reshape(-1, 4), d["ray_d"].reshape(-1, 3)
    if tracer == "host":
        ph = d["photons9"].reshape(-1, 9)
        pos, pw, dr = ph[:, 0:3], ph[:, 3:6], ph[:, 6:9]
    else:
        # device-traced photons stay on the GPU: trace the same path indices through the ABI
        ctx = bre.Bre(device=0)
        ctx.set_medium_homogeneous(m[0:3], m[3:6], float(m[6]))
        ctx.set_lights([dict(kind=0, pos=(0.0, 0.9, 0.0), I=(5.0, 5.0, 5.0))])
        ctx.set_triangles(*fog_box_mesh())     # what the adapter extracts from the pbrt scene
        ctx.trace_photons(0, n_paths, 5, radius)
        pos, pw, dr, _ = ctx.download_photons()
        ctx.close()
    assert pos.shape[0] > n_paths // 2
    oracle.set_medium_homogeneous(m[0:3], m[3:6], float(m[6]))
    oracle.set_photons(pos, pw, dr, float(d["radius"][0]))
    oracle.build()
    rgb, _ = oracle.query(ro, rd, float(d["n_paths"][0]))
    ref = oracle.film_resolve(rgb, 0, 1.0).astype(np.float64)
    assert ref.max() > 0
    err = np.abs(img - ref)
    assert err.mean() <= 1e-4 * np.abs(ref).mean()
    assert err.max() <= 1e-3 * np.abs(ref).max()


@needs_bin
def test_pbrt_adapter_resumes_from_a_checkpoint(tmp_path):
    """"startiteration" / "enditeration" (integrators/photonbeam.cpp:354-358,594-595) + "checkpoint": iterations 0-1 in
    one process, 2-3 in a second one that reloads the film accumulators -- the image equals the uninterrupted 4-iteration
    render (float atomics: within the radiance tolerance).  "imagewritefrequency" keeps the reference's default
    (1 << 31: only the last iteration writes) and "rendersurfaces false" is accepted."""
    def run(out, *extra):
        subprocess.run([BIN, str(tmp_path / out), str(tmp_path / "dump.bin"), "device", "32", "5000", "0.2", "4", "device",
                        "rendersurfaces=false", *extra], check=True, timeout=300)
        return read_pfm(str(tmp_path / out))
    full = run("full.pfm")
    ck = str(tmp_path / "film.ck")
    half = run("half.pfm", "enditeration=2", f"checkpoint={ck}")
    assert os.path.exists(ck) and np.abs(half - full).max() > 0          # two iterations are a different estimate
    resumed = run("resumed.pfm", "startiteration=2", f"checkpoint={ck}")
    err = np.abs(resumed.astype(np.float64) - full)
    assert err.mean() <= 1e-4 * np.abs(full).mean() and err.max() <= 1e-3 * np.abs(full).max()
    # without the checkpoint the reference's own behaviour: L = Ld / (iter + 1) over the iterations run (half the light)
    bare = run("bare.pfm", "startiteration=2")
    assert 0.3 < bare.mean() / full.mean() < 0.7
