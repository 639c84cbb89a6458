#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from the UNMODIFIED reference (needs /root/reference).

    python tests/golden/make_golden.py

Builds oracle/_ref/ref_harness (oracle/Makefile `ref`: the reference sources compiled where they
lie under /root/reference/src) and stores its outputs -- values computed by REFERENCE code -- as
compressed .npz fixtures.  The fixtures travel to the GPU box; /root/reference does not.

  ref_primitives.npz         known-answer vectors per primitive ("<section>/<name>")
  ref_fogbox.npz             photons dumped from the reference's PhotonIntegrator::ShootPhotons on the
                             fog-box scene + reference camera rays + brute-force
                             beam radiance estimate evaluated with reference arithmetic classes
  ref_fogbox_100k.npz        the same at configs[0] size: 100 000 reference photons, 256 x 256 reference camera rays; per-ray
                             hit COUNTS and order-independent set digests (sum of splitmix64(index)) instead of lists
  ref_gridspot.npz           query fixture for a GridDensityMedium lit by a SpotLight: photons walked by reference classes,
                             reference camera rays, brute-force estimate with transmittance from a fine integral of the
                             reference's Density()
  ref_halton.npz             film positions of the camera samples the reference's HaltonSampler hands to the camera pass
                             (StartPixel / SetSampleNumber(iter) / GetCameraSample) for several pixel windows and iterations
  ref_trace_homogeneous.npz  volume-photon deposit loop executed by reference classes
  ref_trace_grid.npz         (sampler stream of photonbeam.cpp:386-391, deposit rule photonmap.cpp:696-724)
  ref_gridtr.npz             GridDensityMedium::Tr statistics + fine integral of the reference's Density() through the
                             64^3 / 256^3 smoke grids of bench.py: pins the optical-depth table at the product's steps
  ref_surfaces.npz           photon paths bouncing off matte triangle walls (photonbeam.cpp:296-323) and
                             crossing an interface-only medium boundary, walked by reference classes;
                             camera rays of a cropped PerspectiveCamera with first-hit tMax
"""
import json
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
DTYPES = {"f4": np.float32, "u4": np.uint32, "u8": np.uint64, "i4": np.int32}


def read_bref(path):
    """Container written by oracle/ref_harness.cpp BinWriter."""
    out = {}
    with open(path, "rb") as f:
        assert f.read(6) == b"BREF1\n"
        while True:
            head = f.read(40)
            if len(head) < 40:
                break
            name = head[:24].rstrip(b"\0").decode()
            dt = DTYPES[head[24:32].rstrip(b"\0").decode()]
            (count,) = struct.unpack("<Q", head[32:40])
            out[name] = np.frombuffer(f.read(count * np.dtype(dt).itemsize), dtype=dt).copy()
    return out


def gridtr(tmp, n_rays=24, n_samples=4_000_000):
    """ref_gridtr.npz: the reference's GridDensityMedium::Tr (mean / standard error of n_samples evaluations) and the
    fine integral of the reference's own Density() along camera rays of bench.py's smoke scenes -- the 64^3 grid of
    smoke10m_1080p and the 256^3 grid of smoke50m_4k -- from the ray origin to a random depth t.  Pins the product's
    deterministic optical-depth table AT THE STEP COUNT IT RUNS WITH (tests/test_oracle_golden.py).  The density
    grids are not stored: bench.smoke_density(n) regenerates them (a checksum is kept)."""
    sys.path.insert(0, ROOT)
    import bench
    out = {}
    for n in (64, 256):
        dens = bench.smoke_density(n)
        dp = os.path.join(tmp, f"dens{n}.bin")
        dens.tofile(dp)
        rng = np.random.default_rng(1000 + n)
        ro, rd = bench.camera_rays(1920, 1080)
        live = np.nonzero(ro[:, 3] > 0)[0]
        pick = rng.choice(live, n_rays, replace=False)
        rays = np.zeros((n_rays, 8), np.float32)
        rays[:, :3] = ro[pick, :3]
        rays[:, 3:6] = rd[pick]
        # the camera sits outside the box: the medium spans roughly t in [tMax - 1.1, tMax]
        rays[:, 6] = ro[pick, 3] - (rng.random(n_rays) * 0.9).astype(np.float32)
        rays[: n_rays // 3, 6] = ro[pick[: n_rays // 3], 3]          # a third of them all the way through
        rp = os.path.join(tmp, f"rays{n}.bin")
        rays.tofile(rp)
        ob = os.path.join(tmp, f"gridtr{n}.bin")
        subprocess.check_call([HARNESS, "gridtr", ob, dp, str(n), str(n), str(n), rp, str(n_samples),
                               str(bench.SIGMA_A[0]), str(bench.SIGMA_S[0]), repr(float(bench.ASPECT)),
                               repr(float(bench.BOX_LO[0]))])
        for k, v in read_bref(ob).items():
            out[f"g{n}/{k}"] = v
        out[f"g{n}/density_sum"] = np.array([dens.astype(np.float64).sum()])
    np.savez_compressed(os.path.join(HERE, "ref_gridtr.npz"), **out)


def fogbox_big(tmp):
    fb = os.path.join(tmp, "fogbox_100k.bin")
    subprocess.check_call([HARNESS, "fogbox", fb, "100000", "256", "0.05", "100000", "digest"], stdout=subprocess.DEVNULL)
    np.savez_compressed(os.path.join(HERE, "ref_fogbox_100k.npz"), **read_bref(fb))


def gridspot(tmp):
    gb = os.path.join(tmp, "gridspot.bin")
    subprocess.check_call([HARNESS, "gridspot", gb, "5000", "40", "0.11"])
    np.savez_compressed(os.path.join(HERE, "ref_gridspot.npz"), **read_bref(gb))


def halton(tmp):
    hb = os.path.join(tmp, "halton.bin")
    subprocess.check_call([HARNESS, "halton", hb])
    np.savez_compressed(os.path.join(HERE, "ref_halton.npz"), **read_bref(hb))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "halton":
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-j8"])
        halton(tempfile.mkdtemp())
        return
    if len(sys.argv) > 1 and sys.argv[1] in ("fogbox_100k", "gridspot"):       # only this fixture
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-j8"])
        {"fogbox_100k": fogbox_big, "gridspot": gridspot}[sys.argv[1]](tempfile.mkdtemp())
        return
    if len(sys.argv) > 1 and sys.argv[1] == "gridtr":       # only this fixture
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-j8"])
        gridtr(tempfile.mkdtemp())
        return
    if not os.path.isdir("/root/reference/src"):
        sys.exit("the reference is not mounted here; fixtures can only be regenerated where it is")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-j8"])
    tmp = tempfile.mkdtemp()
    # primitives
    pj = os.path.join(tmp, "prim.json")
    subprocess.check_call([HARNESS, "primitives", pj])
    flat = {}
    for sec, d in json.load(open(pj)).items():
        for k, v in d.items():
            kind, name = k.split(":")
            a = np.asarray(v, dtype=np.uint32).view(np.float32) if kind == "f" else np.asarray(v, dtype=np.int64)
            flat[f"{sec}/{name}"] = a
    np.savez_compressed(os.path.join(HERE, "ref_primitives.npz"), **flat)
    # fog box: keep the first 8000 photons of the reference dump (small fixture)
    fb = os.path.join(tmp, "fogbox.bin")
    subprocess.check_call([HARNESS, "fogbox", fb, "4000", "32", "0.2", "8000"], stdout=subprocess.DEVNULL)
    np.savez_compressed(os.path.join(HERE, "ref_fogbox.npz"), **read_bref(fb))
    # composed deposit loop
    for kind, first, n in (("homogeneous", 0, 3000), ("grid", 100, 3000)):
        tb = os.path.join(tmp, f"trace_{kind}.bin")
        subprocess.check_call([HARNESS, "trace", tb, kind, str(first), str(n), "5"])
        np.savez_compressed(os.path.join(HERE, f"ref_trace_{kind}.npz"), **read_bref(tb))
    # surfaces: matte walls, reversed wall, interface-only boundary + cropped camera window
    sb = os.path.join(tmp, "surfaces.bin")
    subprocess.check_call([HARNESS, "surfaces", sb, "0", "4000", "6", "64", "48"])
    np.savez_compressed(os.path.join(HERE, "ref_surfaces.npz"), **read_bref(sb))
    gridtr(tmp)
    fogbox_big(tmp)
    gridspot(tmp)
    halton(tmp)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
