#!/usr/bin/env python
"""bench.py -- beam-radiance-estimate hot path on N B200s (one process per GPU).

One "step" = one full iteration of the hot path over one batch of synthetic input:
  K1 trace volume photons (this rank's slice of path indices) -> [N>1: fused push over NVLink / NCCL all-gather]
  -> K2-K4 LBVH build -> K6 optical-depth table + K5 beam query -> K7 film accumulation
  [-> N>1: film gather of the ranks' tiles into rank 0's image].
Workload (BASELINE.json configs[2], the 10M-photon config the roofline target is quoted on):
heterogeneous GridDensityMedium smoke volume (64^3 grid of 8 seeded Gaussian blobs), ~10M volume
photons, 1920x1080 camera rays, fixed gather radius.

Multi-GPU (`--scaling strong`, default): ONE fixed frame on N GPUs.  `--shard photons` (default): the
photon paths are split by index and stay where they were traced -- every rank builds the LBVH of its 1/N
of the map, answers all rays from it, and the partial films are summed into rank 0's image
(bre_set_photon_sharding; the estimate is a sum over photons, integrators/photonbeam.cpp:494-508).
`--shard tiles`: photons all-gathered, the frame split by 16x16 image tile (bre_tile_owner,
integrators/photonbeam.cpp:345-347), film tiles gathered on rank 0.  value = rays of the frame / step time.  `--scaling weak`: every rank renders the whole frame at its own sample offset
(N x samples per pixel), value = N x rays / step time.

`value`   = beam queries/s, whole job, inputs resident in HBM (device-timed, CUDA events).
`e2e`     = the same step through the calls the pbrt adapter makes, host in / host out:
            bre_set_camera (the camera matrices are the step's input) -> photon pass -> build ->
            bre_query_camera_halton_to_film (K0 generates the rays on the device, at the reference's Halton camera
            samples; --camera-samples uniform: bre_query_camera_to_film) -> bre_film_gather to a
            pinned host image.  `e2e_host_rays` (N = 1): the variant that uploads 58 MB of host rays.
`roofline`= K5 (k_tile_walk + k_leaf_tile): DRAM bytes of one launch (ncu, profiles/k5_traffic.json,
            which names the build it was captured on) / kernel time, against the measured HBM copy
            bandwidth.  The SURVEY 8d algorithmic bytes of the canonical per-ray walk are reported
            beside it (`algorithmic_*`): the tile path does not perform that walk.
`cpu_baseline` = CPU figures on the box's host cores for a bounded sample of the same workload.

--impl reference: the UNMODIFIED reference `pbrt` (oracle/_ref/pbrt, built from /root/reference by
oracle/Makefile) rendering the fog box of SURVEY Appendix A with Integrator "photonbeam" on all host
threads, one bounded run per step; queries/s derived as BASELINE.md section 2.  Falls back to the
oracle port when the binary is not there.
"""
import argparse
import json
import math
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

WORKLOADS = {
    # name: (target photons, width, height, expected hits/ray knob h, grid n)
    "smoke10m_1080p": dict(photons=10_000_000, w=1920, h=1080, hits=64, grid=64),
    "smoke1m_540p": dict(photons=1_000_000, w=960, h=540, hits=64, grid=32),
    # BASELINE configs[1]: spotlight through homogeneous fog, 1 M photons, 1080p, fixed radius
    "fog1m_1080p": dict(photons=1_000_000, w=1920, h=1080, hits=64, grid=0),
    # BASELINE configs[3]: 4K camera against 50 M photons (all-gathered over NVLink for N > 1)
    "smoke50m_4k": dict(photons=50_000_000, w=3840, h=2160, hits=64, grid=256),
    "smoke100k_128": dict(photons=100_000, w=128, h=72, hits=32, grid=16),
}
SIGMA_A, SIGMA_S, G = [0.4] * 3, [3.6] * 3, 0.3   # sigma_t = 4 (SURVEY 8d S-smoke)
ASPECT = 16.0 / 9.0
BOX_LO = np.array([0.5 - ASPECT / 2, 0.0, 0.0], np.float32)
BOX_HI = np.array([0.5 + ASPECT / 2, 1.0, 1.0], np.float32)
MAX_DEPTH = 5
EYE = (0.5, 0.5, -1.5)
FOV_DEG = 40.0
PBRT_BIN = os.path.join(ROOT, "oracle", "_ref", "pbrt")
HARNESS_BIN = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def smoke_density(n, seed=8):
    """Sum of 8 seeded Gaussian blobs on an n^3 grid (z,y,x order), plus a thin haze."""
    rng = np.random.default_rng(seed)
    z, y, x = np.meshgrid(*(3 * [(np.arange(n) + 0.5) / n]), indexing="ij")
    d = np.full((n, n, n), 0.02, np.float64)
    for _ in range(8):
        c = rng.uniform(0.2, 0.8, 3)
        s = rng.uniform(0.08, 0.2)
        d += rng.uniform(0.5, 1.0) * np.exp(-((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) / (2 * s * s))
    return d.astype(np.float32)


def world_to_medium():
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = 1.0 / ASPECT
    m[0, 3] = -float(BOX_LO[0]) / ASPECT
    return m


def camera_rays(w, h, row0=0, row1=None, off=(0.5, 0.5)):
    """Pinhole at (0.5,0.5,-1.5) looking +z, vertical fov 40 deg, sample position `off` inside the pixel
    (pixel centres); tMax = exit of the medium box.  float32 (n,4) {o,tMax}, (n,3) unit directions."""
    row1 = h if row1 is None else row1
    eye = np.array(EYE, np.float32)
    tan_half = np.float32(math.tan(math.radians(FOV_DEG / 2)))
    xs = (np.arange(w, dtype=np.float32) + np.float32(off[0])) / np.float32(w)
    ys = (np.arange(row0, row1, dtype=np.float32) + np.float32(off[1])) / np.float32(h)
    sx = (2 * xs - 1) * tan_half * np.float32(w / h)
    sy = (1 - 2 * ys) * tan_half
    d = np.empty((row1 - row0, w, 3), np.float32)
    d[..., 0] = sx[None, :]
    d[..., 1] = sy[:, None]
    d[..., 2] = 1
    d /= np.linalg.norm(d, axis=2, keepdims=True).astype(np.float32)
    d = d.reshape(-1, 3)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = np.float32(1) / d
        t0 = (BOX_LO - eye) * inv
        t1 = (BOX_HI - eye) * inv
    tfar = np.minimum.reduce(np.maximum(t0, t1), axis=1)
    tnear = np.maximum.reduce(np.minimum(t0, t1), axis=1)
    tmax = np.where(tfar > np.maximum(tnear, 0), tfar, 0).astype(np.float32)
    o = np.empty((d.shape[0], 4), np.float32)
    o[:, :3] = eye
    o[:, 3] = tmax
    return o, np.ascontiguousarray(d)


def camera_matrices(w, h):
    """The same pinhole as bre_camera (K0): raster -> camera is affine in (x, y) on the plane z = 1
    (cameras/perspective.cpp:95-103 normalises the camera-space point), camera -> world a translation."""
    th = math.tan(math.radians(FOV_DEG / 2))
    a = w / h
    r2c = np.array([[2 * a * th / w, 0, 0, -a * th], [0, -2 * th / h, 0, th], [0, 0, 0, 1], [0, 0, 0, 1]], np.float32)
    c2w = np.eye(4, dtype=np.float32)
    c2w[:3, 3] = EYE
    return r2c, c2w


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i] == "Active" for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.samples[0][1]) if self.samples[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.samples)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def scene_lights(wl):
    spot = dict(kind=1, pos=(0.5, 0.98, 0.5), I=(30, 30, 30), axis=(0, -1, 0),
                cos_total_width=math.cos(math.radians(35)), cos_falloff_start=math.cos(math.radians(25)))
    if not wl["grid"]:
        return [spot]                       # homogeneous fog: the spotlight alone (configs[1])
    return [dict(kind=0, pos=(0.5, 0.5, 0.5), I=(8, 8, 8)), spot]


def set_medium(obj, wl):
    """Same call on the library context and on the oracle; returns the density grid (None = homogeneous)."""
    if not wl["grid"]:
        obj.set_medium_homogeneous(SIGMA_A, SIGMA_S, G)
        return None
    dens = smoke_density(wl["grid"])
    obj.set_medium_grid(SIGMA_A, SIGMA_S, G, dens, world_to_medium())
    return dens


def setup_scene(ctx, wl):
    dens = set_medium(ctx, wl)
    ctx.set_lights(scene_lights(wl))
    ctx.set_bounds(BOX_LO, BOX_HI)
    return dens


def workload_radius(wl):
    vol = float(np.prod(BOX_HI - BOX_LO))
    return math.sqrt(wl["hits"] * vol / (math.pi * wl["photons"]))


def base_config(name, wl):
    """The keys both arms print (the driver compares them)."""
    return {"workload": name, "photons": wl["photons"], "rays": wl["w"] * wl["h"], "resolution": [wl["w"], wl["h"]],
            "radius": workload_radius(wl), "grid": wl["grid"], "max_depth": MAX_DEPTH,
            "l2": "inputs larger than L2 (photon SoA + tree >> 126 MB at 10M)",
            "step": "K1 trace + [exchange] + K2-4 build + K6 tau + K5 query + K7 film [+ film gather]"}


def calibrate_paths(ctx, wl, radius):
    """Photons stored per emitted path (the workload is defined by its photon count)."""
    ctx.clear_photons()
    n = ctx.trace_photons(0, 200_000, MAX_DEPTH, radius)
    ctx.clear_photons()
    return max(1, int(math.ceil(wl["photons"] / (n / 200_000.0))))


def count_launches(fn):
    """Kernels launched by one call of fn(), counted by CUPTI through torch.profiler (outside the timed region)."""
    try:
        import torch
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        names = [e.name for e in prof.events() if e.device_type.name == "CUDA" and
                 not e.name.lower().startswith(("memcpy", "memset"))]
        ours = [n for n in names if "bre::" in n or n.startswith(("k_", "void k_"))]
        return len(ours), len(names)
    except Exception:
        return None, None


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bre = ge.load_bre()
    wl = WORKLOADS[args.workload]
    uid = None
    if world > 1:
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            import ctypes
            raw = ctypes.create_string_buffer(128)
            assert bre.load_library().bre_nccl_unique_id(raw) == 0
            buf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).cuda()
        dist.broadcast(buf, 0)
        uid = bytes(buf.cpu().numpy().tobytes())
    ctx = bre.Bre(device=local, n_ranks=world, rank=rank, nccl_unique_id=uid, max_leaf_photons=args.leaf)
    setup_scene(ctx, wl)
    radius = workload_radius(wl)
    n_paths = calibrate_paths(ctx, wl, radius)       # identical on every rank (deterministic)
    strong = args.scaling == "strong"
    p_first, p_count = bre.shard_range(n_paths, world, rank, ctx.lib)
    W, H = wl["w"], wl["h"]
    # weak scaling: rank r renders the whole frame at its own sample position inside the pixel
    off = (0.5, 0.5) if (strong or world == 1) else ((rank + 0.5) / world, ((rank * 7) % world + 0.5) / world)
    ro_h, rd_h = camera_rays(W, H, off=off)
    nq = ro_h.shape[0]
    img_pin = torch.empty((nq, 3), dtype=torch.float32).pin_memory()
    ro_d = torch.from_numpy(ro_h).cuda()
    rd4 = np.zeros((nq, 4), np.float32); rd4[:, :3] = rd_h
    rd_d = torch.from_numpy(rd4).cuda()
    out_d = torch.empty((nq, 4), dtype=torch.float32, device="cuda")
    ctx.film_create(nq)
    ctx.set_timing(True)
    tile = not args.no_tile
    if tile:
        ctx.set_ray_grid(W, H)          # the rays are a W x H image: K5 walks the LBVH once per 16x16 tile
    sharded = strong and world > 1
    if sharded:
        if not tile:
            raise SystemExit("--scaling strong on several GPUs needs the tile path")
        if args.shard == "photons":
            ctx.set_photon_sharding(True)   # every rank keeps its photons: 1/N trace, build and photon tests, films summed
        else:
            ctx.set_tile_sharding(True)     # photons replicated, 1/N of the image tiles per rank
    by_photon = sharded and args.shard == "photons"
    r2c, c2w = camera_matrices(W, H)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))

    def photon_pass():
        # N > 1: fused trace + push over NVLink P2P (--exchange p2p, default) or trace + NCCL all-gather
        if world > 1 and (args.exchange == "p2p" or by_photon):
            ctx.trace_exchange_photons(n_paths, MAX_DEPTH, radius)   # photon sharding: the slice stays local
            return
        ctx.clear_photons()
        ctx.trace_photons(p_first, p_count, MAX_DEPTH, radius)
        if world > 1:
            ctx.allgather_photons()

    def step_device():
        # the rays of the pass are known before the photons: their optical-depth tables (K6) run on the staging
        # stream under the photon pass and the build (bre_stage_rays_d), like the adapter's bre_stage_camera
        if not args.no_stage:
            ctx.stage_rays_d(nq, ro_d.data_ptr(), rd_d.data_ptr())
        photon_pass()
        ctx.build()
        ctx.query_d(nq, ro_d.data_ptr(), rd_d.data_ptr(), float(n_paths), out_d.data_ptr())
        ctx.film_accumulate_d(nq, out_d.data_ptr())
        if sharded:
            ctx.film_gather_ptr(0, 1.0, None)           # tiles -> rank 0's image (stays on the device)

    def step_e2e():
        # the calls the pbrt adapter makes: camera in, image out
        ctx.set_camera(r2c, c2w, 0, 0, W, H, True)
        halton = args.camera_samples == "halton" and (strong or world == 1)
        if not args.no_stage:
            if halton:
                ctx.stage_camera_halton(0)              # K0 (the reference's per-pixel Halton camera samples) + K6 under the photon pass
            else:
                ctx.stage_camera(off[0], off[1])        # one sub-pixel offset for every pixel
        photon_pass()
        ctx.build()
        if halton:
            ctx.query_camera_halton_to_film(float(n_paths), 0)
        else:
            ctx.query_camera_to_film(float(n_paths), off[0], off[1])
        if sharded:
            ctx.film_gather_ptr(0, 1.0, img_pin.data_ptr() if rank == 0 else None)
        else:
            ctx.film_resolve_ptr(0, 1.0, img_pin.data_ptr())

    ro_pin = rd_pin = None

    def step_e2e_host_rays():
        ctx.stage_rays_ptr(nq, ro_pin.data_ptr(), rd_pin.data_ptr())
        photon_pass()
        ctx.build()
        ctx.query_staged_to_film(float(n_paths))
        ctx.film_resolve_ptr(0, 1.0, img_pin.data_ptr())

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, steps, per_step=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
            if per_step:
                per_step()
        e1.record(stream)
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        dev = e0.elapsed_time(e1)
        ms = max(dev, wall)   # host-side syncs inside a step are part of the step
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        step_device()
    n_photons = ctx.photon_count()
    sampler = ClockSampler(local)
    sampler.start()
    phases = {k: [] for k in ("trace", "build", "tau", "query", "query_a", "query_b", "film")}

    def grab():
        t = ctx.timings()
        for k in phases:
            phases[k].append(t[k])

    ms_total = timed(step_device, args.steps, per_step=grab)
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    ms_e2e_rays = None
    if world == 1 and not args.no_host_rays:
        ro_pin = torch.from_numpy(ro_h).pin_memory()
        rd_pin = torch.from_numpy(rd_h).pin_memory()
        step_e2e_host_rays()
        ms_e2e_rays = timed(step_e2e_host_rays, max(2, args.steps // 2)) / max(2, args.steps // 2)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    ms_step = ms_total / args.steps
    total_rays = nq if strong else world * nq
    value = total_rays / (ms_step * 1e-3)
    e2e_value = total_rays / (ms_e2e / args.steps * 1e-3)
    ours_launches, all_launches = count_launches(step_device)

    # canonical counters of one K5 launch (outside the timed region), on rank 0's view of the whole frame
    ctx.set_tile_sharding(False)
    cnt = ctx.count_d(nq, ro_d.data_ptr(), rd_d.data_ptr())
    ph = {k: float(np.mean(v)) for k, v in phases.items()}
    k5_ms = ph["query"]
    peak, peak_src = measured_peak()
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "k5_traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            traffic, traffic_src = tj.get(args.workload), tj.get("_source")
        except Exception:
            traffic = None
    own_bytes = 48 * nq + 64 * cnt["fast_nodes"] + 16 * cnt["fast_disk_tests"] + 32 * cnt["hits"] + 8 * 2 * cnt["leaf_pairs"]
    dram_bytes = traffic if (traffic and world == 1 and tile) else None
    achieved = (dram_bytes if dram_bytes else own_bytes) / (k5_ms * 1e-3) / 1e9

    if rank != 0:
        ctx.close()
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1 and not args.no_cpu:
        cpu = cpu_baseline(ctx, wl, ro_h, rd_h, float(n_paths), args.cpu_seconds)
    line = {
        "metric": "beam_queries_per_s", "value": value, "unit": "queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": args.scaling if world > 1 else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": base_config(args.workload, wl),
        "detail": {"photons_traced": int(n_photons), "photon_paths": int(n_paths), "rays_per_step": int(total_rays),
                   "max_leaf_photons": ctx_leaf(args), "k5_path": "image tiles" if tile else "ray packets",
                   "multi_gpu": ((("photon paths sharded: per-rank trace + LBVH + query of 1/N of the map, films summed on rank 0"
                                   if by_photon else "16x16 image tiles round-robin over ranks + film gather to rank 0")
                                  if strong else "every rank renders the whole frame at its own sample offset")
                                 if world > 1 else None),
                   "exchange": (args.exchange if world > 1 else None),
                   "mean_hits_per_ray": cnt["hits"] / nq, "mean_V": cnt["nodes_visited"] / nq,
                   "mean_T": cnt["disk_tests"] / nq, "mdisk_tests_per_s": cnt["fast_disk_tests"] / (k5_ms * 1e-3) / 1e6,
                   "phase_ms": {"trace": ph["trace"], "build": ph["build"], "tau": ph["tau"], "query": k5_ms,
                                "query_k5a_walk": ph["query_a"], "query_k5b_leaf": ph["query_b"], "film": ph["film"]},
                   "fast_nodes_per_ray": cnt["fast_nodes"] / nq, "fast_disk_tests_per_ray": cnt["fast_disk_tests"] / nq,
                   "leaf_pairs_per_ray": cnt["leaf_pairs"] / nq, "pair_queue_overflow": cnt["pair_queue_overflow"]},
        "e2e": {"value": e2e_value, "unit": "queries/s", "h2d_bytes_per_step": 152,
                "d2h_bytes_per_step": int(nq * 12), "ms_per_step": ms_e2e / args.steps,
                "camera_samples": args.camera_samples,
                "api": ("bre_set_camera + bre_stage_camera_halton + photon pass + bre_build + bre_query_camera_halton_to_film + bre_film_gather/resolve"
                        if args.camera_samples == "halton" else
                        "bre_set_camera + bre_stage_camera + photon pass + bre_build + bre_query_camera_to_film + bre_film_gather/resolve")},
        "gpu_launches": (ours_launches * args.steps) if ours_launches is not None else None,
        "gpu_launches_per_step": ours_launches, "gpu_launches_per_step_all": all_launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": dram_bytes,
                     "traffic_source": (traffic_src if dram_bytes else
                                        "no ncu capture for this configuration: bytes of the production path's own structures "
                                        "(rays + 64 B nodes + 16 B photon tests + 32 B hits + 16 B pairs), an upper bound of its DRAM traffic"),
                     "kernel": ("K5 = k_tile_walk<0,false> + k_leaf_tile<0,false,8> (+ k_traverse / k_leaf_pairs for incoherent tiles)"
                                if tile else "K5 = k_traverse<0,false> + k_leaf_fat<0,false,8> + k_leaf_pairs<0,false,8>"),
                     "kernel_ms": k5_ms, "own_structure_bytes": int(own_bytes),
                     "algorithmic_bytes": int(cnt["bytes_algorithmic"]),
                     "algorithmic_gbs": cnt["bytes_algorithmic"] / (k5_ms * 1e-3) / 1e9,
                     "algorithmic_speedup": cnt["bytes_algorithmic"] / (k5_ms * 1e-3) / 1e9 / peak,
                     "note": "frac = DRAM bytes / kernel time / peak. algorithmic_* counts SURVEY 8d's canonical per-ray walk "
                             "(48 + 32 V + 16 T + 32 H per ray), which the tile path does not perform: a work ratio, not a roofline fraction",
                     "peak_source": peak_src},
        "clocks": sampler.summary(),
    }
    if ms_e2e_rays is not None:
        line["e2e_host_rays"] = {"value": nq / (ms_e2e_rays * 1e-3), "unit": "queries/s", "ms_per_step": ms_e2e_rays,
                                 "h2d_bytes_per_step": int(nq * 28), "d2h_bytes_per_step": int(nq * 12),
                                 "api": "bre_stage_rays + photon pass + bre_build + bre_query_staged_to_film + bre_film_resolve"}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def ctx_leaf(args):
    return args.leaf if args.leaf > 0 else 8


def cpu_baseline(ctx, wl, ro_h, rd_h, n_paths, seconds):
    """CPU figures on a bounded sample of the same workload, all host cores: the oracle port on the same
    photons and rays (kind "port"), and -- when oracle/_ref/ref_harness is there -- the reference's own
    PhotonBeamBVH code on a subset of the same photons (reference_code)."""
    orc = ge.load_orc()
    o = orc.Oracle()
    pos, pw, dr, rad = ctx.download_photons()
    out = _cpu_run(o, pos, pw, dr, rad, ro_h, rd_h, n_paths, seconds, wl)
    rc = reference_code_query(pos, rad, ro_h, rd_h, min(seconds, 8.0))
    if rc:
        out["reference_code"] = rc
    return out


def reference_code_query(pos, rad, ro_h, rd_h, seconds, max_photons=1_000_000, max_rays=65536):
    """the reference's PhotonBeamBVH (SAH build + Intersect) on a subset of the photons."""
    if not os.path.exists(HARNESS_BIN):
        return None
    rng = np.random.default_rng(2)
    n = pos.shape[0]
    keep = rng.choice(n, min(n, max_photons), replace=False) if n > max_photons else np.arange(n)
    ph = np.empty((keep.size, 4), np.float32)
    ph[:, :3] = pos[keep]
    ph[:, 3] = rad[keep] * np.float32(math.sqrt(n / keep.size))     # same expected hits per ray as the full map
    qi = rng.integers(0, ro_h.shape[0], min(max_rays, ro_h.shape[0]))
    ry = np.zeros((qi.size, 8), np.float32)
    ry[:, :3] = ro_h[qi, :3]; ry[:, 3:6] = rd_h[qi]; ry[:, 6] = ro_h[qi, 3]
    cores = os.cpu_count() or 1
    with tempfile.TemporaryDirectory() as tmp:
        pp, rp = os.path.join(tmp, "ph.bin"), os.path.join(tmp, "ry.bin")
        ph.tofile(pp); ry.tofile(rp)
        try:
            r = subprocess.run([HARNESS_BIN, "bvhquery", pp, rp, str(cores), str(seconds)], capture_output=True,
                               text=True, timeout=300)
            d = json.loads(r.stdout.strip().splitlines()[-1])
        except Exception as e:       # the harness is test infrastructure: its absence must not fail the bench
            return {"error": str(e)[:200]}
    return {"value": d["queries_per_s"], "unit": "queries/s", "cores": cores, "kind": "reference",
            "sample": f"reference PhotonBeamBVH (SAH build {d['build_s']:.1f} s, 1 thread) over {d['photons']} of the "
                      f"{n} photons as short beams, PhotonBeamBVH::Intersect + disk predicate for {d['rays']} camera rays "
                      f"on {cores} threads ({d['query_s']:.1f} s); {d['candidates'] / max(1, d['rays']):.0f} candidates/ray"}


def _cpu_run(o, pos, pw, dr, rad, ro_h, rd_h, n_paths, seconds, wl):
    set_medium(o, wl)
    o.set_photons(pos, pw, dr, rad)
    t0 = time.perf_counter()
    o.build()
    t_build = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(1)
    nq = ro_h.shape[0]
    done, spent, batch = 0, 0.0, 4096
    while spent < seconds and done < nq:
        idx = rng.integers(0, nq, batch)
        t0 = time.perf_counter()
        o.query(ro_h[idx], rd_h[idx], n_paths, threads=cores)
        dt = time.perf_counter() - t0
        spent += dt
        done += batch
        if dt < 1.0:
            batch = min(batch * 2, 1 << 18)
    return {"value": done / spent, "unit": "queries/s", "cores": cores, "kind": "port",
            "sample": f"{done} random rays of the same {nq}-ray camera against the same {pos.shape[0]} photons, "
                      f"query only ({spent:.1f} s; CPU tree build {t_build:.1f} s excluded)"}


# ---------------------------------------------------------------------------------------------- reference arm
FOG_SCENE = """LookAt 0 0 -3.5  0 0 0  0 1 0
Camera "perspective" "float fov" [40]
Film "image" "integer xresolution" [256] "integer yresolution" [256] "string filename" ["{out}"] "float cropwindow" [{c0} {c1} {c0} {c1}]
Sampler "halton" "integer pixelsamples" [1]
Integrator "photonbeam" "integer iterations" [1] "integer photonsperiteration" [{photons}] "float initialbeamradius" [0.05] "integer maxdepth" [5]
MakeNamedMedium "fog" "string type" "homogeneous" "rgb sigma_a" [0.05 0.05 0.05] "rgb sigma_s" [0.4 0.4 0.4] "float g" [0.0]
MediumInterface "" "fog"
WorldBegin
AttributeBegin
  MediumInterface "fog" "fog"
  LightSource "point" "point from" [0 0.9 0] "rgb I" [5 5 5]
AttributeEnd
AttributeBegin
  MediumInterface "fog" ""
  Material "matte" "rgb Kd" [0.6 0.6 0.6]
  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [-4 -1 -4  4 -1 -4  4 -1 4  -4 -1 4]
  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [-4 2 -4  -4 2 4  4 2 4  4 2 -4]
  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [-4 -1 4  4 -1 4  4 2 4  -4 2 4]
  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [-4 -1 -4  -4 2 -4  4 2 -4  4 -1 -4]
  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [-4 -1 -4  -4 -1 4  -4 2 4  -4 2 -4]
  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [4 -1 -4  4 2 -4  4 2 4  4 -1 4]
AttributeEnd
WorldEnd
"""


def run_pbrt_once(tmp, photons, crop, threads):
    """One run of the unmodified reference: pbrt --nthreads T scene.pbrt with Integrator "photonbeam"
    (core/api.cpp:1463-1464, main/pbrt.cpp:82-86).  Returns wall seconds, camera-segment queries
    (photon-surface interactions minus beams, integrators/photonbeam.cpp:290,489, as BASELINE.md section 2),
    pbrt's own profile split and the core count of its banner (main/pbrt.cpp:134-135)."""
    scene = os.path.join(tmp, "fog.pbrt")
    c0, c1 = 0.5 - crop / 2, 0.5 + crop / 2
    with open(scene, "w") as f:
        f.write(FOG_SCENE.format(out=os.path.join(tmp, "out.pfm"), c0=c0, c1=c1, photons=photons))
    t0 = time.perf_counter()
    r = subprocess.run([PBRT_BIN, "--nthreads", str(threads), scene], capture_output=True, text=True, timeout=1800)
    wall = time.perf_counter() - t0
    txt = (r.stdout + r.stderr).replace("\r", "\n")
    if r.returncode != 0:
        raise RuntimeError("reference pbrt failed: " + txt[-400:])

    def num(pat):
        m = re.search(pat, txt)
        return int(m.group(1)) if m else None

    def secs(name):
        m = re.search(re.escape(name) + r"\s+[\d.]+%\s+\(\s*(\d+):(\d+):([\d.]+)\)", txt)
        return (int(m.group(1)) * 3600 + int(m.group(2)) * 60 + float(m.group(3))) if m else None
    m = re.search(r"PhotonBeamBVH\s*\n\s*Interior nodes\s+\d+\s*\n\s*Leaf nodes\s+(\d+)", txt)
    beams = int(m.group(1)) if m else None
    inter = num(r"Visible points checked per photon intersection\s+\d+ /\s+(\d+)")
    cores = num(r"\[Detected (\d+) cores\]")
    queries = (inter - beams) if (beams is not None and inter is not None) else None
    return {"wall_s": wall, "queries": queries, "beams": beams, "detected_cores": cores,
            "profile_s": {"photon_pass": secs("SPPM photon pass"), "bvh_build": secs("PhotonBeamBVHConstruction"),
                          "camera_pass": secs("SPPM camera pass"), "bvh_query": secs("PhotonBeamBVHQuery")}}


def port_figure(wl, seconds=8.0):
    """The oracle port end to end on the benchmark's own workload (kind "port"): CPU photon pass + tree build once,
    then a bounded sample of the camera rays on all host threads."""
    orc = ge.load_orc()
    o = orc.Oracle()
    set_medium(o, wl)
    lights = scene_lights(wl)
    radius = workload_radius(wl)
    t0 = time.perf_counter()
    probe, _, _, _ = o.trace(lights, BOX_LO, BOX_HI, 0, 200_000, MAX_DEPTH)
    n_paths = int(math.ceil(wl["photons"] / (probe.shape[0] / 200_000.0)))
    pos, pw, dr, _ = o.trace(lights, BOX_LO, BOX_HI, 0, n_paths, MAX_DEPTH)
    t_trace = time.perf_counter() - t0
    ro_h, rd_h = camera_rays(wl["w"], wl["h"])
    rad = np.full(pos.shape[0], radius, np.float32)
    r = _cpu_run(o, pos, pw, dr, rad, ro_h, rd_h, float(n_paths), seconds, wl)
    r["sample"] += f"; CPU photon pass {t_trace:.1f} s (1 thread)"
    return r


def run_reference(args):
    """CPU arm.  Rank 0 alone works; the other ranks exit at once."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    line = {"impl": "reference", "metric": "beam_queries_per_s", "unit": "queries/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": base_config(args.workload, wl)}
    if os.path.exists(PBRT_BIN):
        runs = []
        with tempfile.TemporaryDirectory() as tmp:
            for i in range(args.warmup + args.steps):
                r = run_pbrt_once(tmp, args.ref_photons, args.ref_crop, cores)
                if i >= args.warmup:
                    runs.append(r)
        wall = float(np.mean([r["wall_s"] for r in runs]))
        q = float(np.mean([r["queries"] or 0 for r in runs]))
        value = q / wall
        prof = {k: float(np.mean([r["profile_s"][k] or 0.0 for r in runs])) for k in runs[0]["profile_s"]}
        npx = int(round(256 * args.ref_crop)) ** 2
        sample = (f"UNMODIFIED reference pbrt (oracle/_ref/pbrt), Integrator \"photonbeam\", --nthreads {cores} "
                  f"(banner: {runs[0]['detected_cores']} cores): fog box of SURVEY Appendix A, {args.ref_photons} photon paths "
                  f"-> {runs[0]['beams']} beams, crop of ~{npx} pixels of 256x256, 1 iteration per step; "
                  f"{q:.0f} camera-segment BVH queries in {wall:.2f} s wall per step "
                  f"(pbrt profile: photon pass {prof['photon_pass']:.2f} s serial by design, BVH build {prof['bvh_build']:.2f} s, "
                  f"camera pass {prof['camera_pass']:.2f} s of which PhotonBeamBVHQuery {prof['bvh_query']:.2f} s). "
                  f"The reference cannot run the 10 M-photon benchmark scene in a bounded step (serial photon pass, "
                  f"shared_ptr beam BVH), and its registered integrator gathers beam x beam, not photon disks (SURVEY 0)")
        line.update({"value": value, "ms_per_step": wall * 1e3,
                     "cpu_baseline": {"value": value, "unit": "queries/s", "cores": cores, "kind": "reference",
                                      "sample": sample, "pbrt_profile_s": prof},
                     "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        if not args.no_port:
            try:
                line["cpu_baseline"]["port"] = port_figure(wl)
            except Exception as e:
                line["cpu_baseline"]["port"] = {"error": str(e)[:200]}
    else:
        # the reference binary did not travel: the oracle port, one bounded sample per step
        times, rate = [], None
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            rate = port_figure(wl, seconds=4.0)
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
        line.update({"value": rate["value"], "ms_per_step": float(np.mean(times)) * 1e3,
                     "cpu_baseline": dict(rate, kind="port"),
                     "e2e": {"value": rate["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="smoke10m_1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--leaf", type=int, default=0)
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--camera-samples", default="halton", choices=["halton", "uniform"],
                    help="e2e arm: the reference's per-pixel Halton camera samples (the adapter's default) or one offset for all pixels")
    ap.add_argument("--shard", default="photons", choices=["photons", "tiles"],
                    help="strong scaling on N > 1 GPUs: shard the photon paths (each rank builds and queries its own "
                         "1/N of the map, films summed) or the image tiles (photons replicated)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-tile", action="store_true", help="K5 packet walk instead of the image-tile walk (A/B)")
    ap.add_argument("--no-host-rays", action="store_true", help="skip the second end-to-end figure (host rays)")
    ap.add_argument("--no-port", action="store_true", help="reference arm: skip the oracle-port figure")
    ap.add_argument("--no-stage", action="store_true", help="K0 / K6 inside the query call instead of under the photon pass (A/B)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-photons", type=int, default=50000, help="reference arm: photon paths per pbrt run")
    ap.add_argument("--ref-crop", type=float, default=0.12, help="reference arm: side of the crop window (fraction of 256)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
