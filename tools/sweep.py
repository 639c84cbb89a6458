#!/usr/bin/env python
"""tools/sweep.py -- synthetic beam-query sweep (BASELINE.json configs[4], SURVEY.md 8d "S-uniform").

Isolates the query kernel: photons uniform in the unit cube (numpy generator -- this is a
throughput tool, parity lives in tests/), fixed gather radius chosen for an expected number of hits
per unit-length ray h:  r = sqrt(h / (pi N)),  homogeneous medium, rays either a pinhole camera at
(0.5,0.5,-1.5) / fov 40 deg with tMax = exit of the cube, or incoherent (random origin on the cube
surface, random direction).  For every (N, h, Q, kind) it prints one JSON line with the K2-K4 build
time, the K5a / K5b kernel times (CUDA events inside libbre.so), queries/s, disk tests/s and the
roofline figure of SURVEY.md 8d (algorithmic bytes from the canonical counters / K5 time).

    python tools/sweep.py --photons 1e6 1e7 --hits 8 64 --rays 65536 2073600 [--incoherent] [--reps 5]
"""
import argparse
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def pinhole(q):
    w = int(round(math.sqrt(q * 16 / 9)))
    h = max(1, q // w)
    eye = np.array([0.5, 0.5, -1.5], np.float32)
    th = np.float32(math.tan(math.radians(20.0)))
    xs = (np.arange(w, dtype=np.float32) + 0.5) / w
    ys = (np.arange(h, dtype=np.float32) + 0.5) / h
    d = np.empty((h, w, 3), np.float32)
    d[..., 0] = ((2 * xs - 1) * th * np.float32(w / h))[None, :]
    d[..., 1] = ((1 - 2 * ys) * th)[:, None]
    d[..., 2] = 1
    d /= np.linalg.norm(d, axis=2, keepdims=True).astype(np.float32)
    d = d.reshape(-1, 3)
    o4, dd = clip_to_cube(np.broadcast_to(eye, d.shape).copy(), d)
    pinhole.last_wh = (w, h)
    return o4, dd


def incoherent(q, seed=4):
    rng = np.random.default_rng(seed)
    o = rng.random((q, 3), dtype=np.float32)
    v = rng.normal(size=(q, 3)).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    return clip_to_cube(o, v)


def clip_to_cube(o, d):
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = np.float32(1) / d
        t0, t1 = (0 - o) * inv, (1 - o) * inv
    tfar = np.minimum.reduce(np.maximum(t0, t1), axis=1)
    tnear = np.maximum.reduce(np.minimum(t0, t1), axis=1)
    o4 = np.empty((o.shape[0], 4), np.float32)
    o4[:, :3] = o
    o4[:, 3] = np.where(tfar > np.maximum(tnear, 0), tfar, 0)
    return o4, np.ascontiguousarray(d, np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--photons", type=float, nargs="+", default=[1e6, 1e7])
    ap.add_argument("--hits", type=float, nargs="+", default=[8, 64])
    ap.add_argument("--rays", type=int, nargs="+", default=[65536, 2073600])
    ap.add_argument("--incoherent", action="store_true")
    ap.add_argument("--no-grid", action="store_true", help="pinhole rays take the per-ray / packet walk instead of the tile path")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--leaf", type=int, default=0)
    ap.add_argument("--g", type=float, default=0.3)
    args = ap.parse_args()
    import torch
    bre = ge.load_bre()
    peak = 6550.0
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = float(json.load(open(pk))["hbm_gbs"])
    for n in [int(x) for x in args.photons]:
        rng = np.random.default_rng(1)
        pos = rng.random((n, 3), dtype=np.float32)
        v = rng.normal(size=(n, 3)).astype(np.float32)
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        pw = rng.random((n, 3), dtype=np.float32) / np.float32(n)
        ctx = bre.Bre(device=0, max_leaf_photons=args.leaf)
        ctx.set_medium_homogeneous([0.05] * 3, [0.4] * 3, args.g)
        ctx.set_timing(True)
        ctx.upload_photons(pos, pw, v, radius=1.0)
        del pos, v, pw
        first_build = True
        for h in args.hits:
            r = math.sqrt(h / (math.pi * n))
            ctx.set_radius(r)
            ctx.build()                      # the first one is a full build, later ones refit
            build_kind = "full" if first_build else "refit"
            first_build = False
            build_ms = ctx.timings()["build"]
            for q in args.rays:
                for kind in (["pinhole", "incoherent"] if args.incoherent else ["pinhole"]):
                    ro, rd = pinhole(q) if kind == "pinhole" else incoherent(q)
                    nq = ro.shape[0]
                    # pinhole rays are a w x h image: the image-tile path of K5 (bre_set_ray_grid) unless --no-grid
                    if kind == "pinhole" and not args.no_grid:
                        ctx.set_ray_grid(*pinhole.last_wh)
                    else:
                        ctx.set_ray_grid(0, 0)
                    ro_d = torch.from_numpy(ro).cuda()
                    rd4 = np.zeros((nq, 4), np.float32); rd4[:, :3] = rd
                    rd_d = torch.from_numpy(rd4).cuda()
                    out_d = torch.empty((nq, 4), dtype=torch.float32, device="cuda")
                    ka, kb = [], []
                    for i in range(args.reps + 2):
                        ctx.query_d(nq, ro_d.data_ptr(), rd_d.data_ptr(), float(n), out_d.data_ptr())
                        t = ctx.timings()
                        if i >= 2:
                            ka.append(t["query_a"]); kb.append(t["query_b"])
                    cnt = ctx.count_d(nq, ro_d.data_ptr(), rd_d.data_ptr())
                    k5 = float(np.median(ka) + np.median(kb))
                    print(json.dumps({
                        "photons": n, "hits_knob": h, "radius": r, "rays": nq, "kind": kind,
                        "k5_path": "tiles" if (kind == "pinhole" and not args.no_grid) else "per-ray",
                        "build": build_kind, "build_ms": round(build_ms, 3),
                        "k5a_ms": round(float(np.median(ka)), 3), "k5b_ms": round(float(np.median(kb)), 3),
                        "mqueries_per_s": round(nq / k5 / 1e3, 2),
                        "gdisk_tests_per_s": round(cnt["fast_disk_tests"] / k5 / 1e6, 2),
                        "hits_per_ray": round(cnt["hits"] / nq, 2), "V": round(cnt["nodes_visited"] / nq, 1),
                        "T": round(cnt["disk_tests"] / nq, 1),
                        "fast_nodes_per_ray": round(cnt["fast_nodes"] / nq, 1),
                        "pairs_per_ray": round(cnt["leaf_pairs"] / nq, 1),
                        "overflow_pairs": cnt["pair_queue_overflow"],
                        "algorithmic_gbs": round(cnt["bytes_algorithmic"] / k5 / 1e6, 1),
                        "roofline_frac": round(cnt["bytes_algorithmic"] / k5 / 1e6 / peak, 3)}), flush=True)
                    del ro_d, rd_d, out_d
        ctx.close()


if __name__ == "__main__":
    main()
