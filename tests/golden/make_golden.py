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
  ref_trace_homogeneous.npz  volume-photon deposit loop executed by reference classes
  ref_trace_grid.npz         (sampler stream of photonbeam.cpp:386-391, deposit rule photonmap.cpp:696-724)
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


def main():
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
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
