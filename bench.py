#!/usr/bin/env python
"""bench.py -- beam-radiance-estimate hot path on N B200s (one process per GPU).

One "step" = one full iteration of the hot path over one batch of synthetic input:
  K1 trace volume photons (this rank's slice of path indices) -> [N>1: NCCL all-gather]
  -> K2-K4 LBVH build -> K6 optical-depth table + K5 beam query over this rank's camera rays
  -> K7 film accumulation.
Workload (BASELINE.json configs[2], the 10M-photon config the roofline target is quoted on):
heterogeneous GridDensityMedium smoke volume (64^3 grid of 8 seeded Gaussian blobs), ~10M volume
photons, 1920x1080 camera rays, fixed gather radius.

`value`   = beam queries/s, whole job, inputs resident in HBM (device-timed, CUDA events).
`e2e`     = the same step through the host C-ABI: pinned host rays -> bre_stage_rays (async H2D) ->
            photon pass -> build -> bre_query_staged_to_film -> bre_film_resolve to a host image;
            host<->device copies inside the timed region.
`roofline`= K5 query kernel: algorithmic bytes (48 + 32 V + 16 T + 32 H per ray,
            V/T/H counted on the canonical 1-photon-per-leaf LBVH) / kernel time, vs the measured
            HBM copy bandwidth in MEASURED_PEAKS.json.
`cpu_baseline` = the CPU oracle (port) on a bounded sample of the same rays, all host cores.

--impl reference: times the CPU implementation alone (oracle/_ref harness when present, else the
oracle port) on the same workload and prints the same JSON shape with "impl": "reference".
"""
import argparse
import json
import math
import os
import subprocess
import sys
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


def camera_rays(w, h, row0=0, row1=None):
    """Pinhole at (0.5,0.5,-1.5) looking +z, vertical fov 40 deg, pixel centres;
    tMax = exit of the medium box.  Returns float32 (n,4) {o,tMax} and (n,3) unit directions."""
    row1 = h if row1 is None else row1
    eye = np.array([0.5, 0.5, -1.5], np.float32)
    tan_half = np.float32(math.tan(math.radians(20.0)))
    xs = (np.arange(w, dtype=np.float32) + np.float32(0.5)) / np.float32(w)
    ys = (np.arange(row0, row1, dtype=np.float32) + np.float32(0.5)) / np.float32(h)
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


def calibrate_paths(ctx, wl, radius):
    """Photons stored per emitted path (the workload is defined by its photon count)."""
    ctx.clear_photons()
    n = ctx.trace_photons(0, 200_000, MAX_DEPTH, radius)
    ctx.clear_photons()
    return max(1, int(math.ceil(wl["photons"] / (n / 200_000.0))))


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
    vol = float(np.prod(BOX_HI - BOX_LO))
    radius = math.sqrt(wl["hits"] * vol / (math.pi * wl["photons"]))
    n_paths = calibrate_paths(ctx, wl, radius)       # identical on every rank (deterministic)
    # weak scaling: every rank owns the full camera (its own spp), photon paths are split by index
    # and all-gathered, so per-GPU query work is fixed as N grows.
    p_first, p_count = bre.shard_range(n_paths, world, rank, ctx.lib)
    W, H = wl["w"], wl["h"]
    ro_h, rd_h = camera_rays(W, H)
    nq = ro_h.shape[0]
    ro_pin = torch.from_numpy(ro_h).pin_memory()
    rd_pin = torch.from_numpy(rd_h).pin_memory()
    img_pin = torch.empty((nq, 3), dtype=torch.float32).pin_memory()
    ro_d = torch.from_numpy(ro_h).cuda()
    rd4 = np.zeros((nq, 4), np.float32); rd4[:, :3] = rd_h
    rd_d = torch.from_numpy(rd4).cuda()
    out_d = torch.empty((nq, 4), dtype=torch.float32, device="cuda")
    ctx.film_create(nq)
    ctx.set_timing(True)
    if not args.no_tile:
        ctx.set_ray_grid(W, H)          # the rays are a W x H image: K5 walks the LBVH once per 16x16 tile
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))

    def photon_pass():
        # N > 1: fused trace + push over NVLink P2P (--exchange p2p, default) or trace + NCCL all-gather
        if world > 1 and args.exchange == "p2p":
            ctx.trace_exchange_photons(n_paths, MAX_DEPTH, radius)
            return
        ctx.clear_photons()
        ctx.trace_photons(p_first, p_count, MAX_DEPTH, radius)
        if world > 1:
            ctx.allgather_photons()

    def step_device():
        photon_pass()
        ctx.build()
        ctx.query_d(nq, ro_d.data_ptr(), rd_d.data_ptr(), float(n_paths), out_d.data_ptr())
        ctx.film_accumulate_d(nq, out_d.data_ptr())

    def step_e2e():
        # the calls the pbrt adapter makes: rays handed over first (async upload from pinned host
        # memory overlaps the photon pass), photon pass, build, camera pass, image back to the host
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

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
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
    # kernel-only timings: collect K5 duration of every timed step via the library's own events
    barrier()
    k5, ktau, kbuild, ktrace, kfilm, k5a, k5b = [], [], [], [], [], [], []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_device()
        t = ctx.timings()
        k5.append(t["query"]); ktau.append(t["tau"]); kbuild.append(t["build"]); ktrace.append(t["trace"]); kfilm.append(t["film"]); k5a.append(t["query_a"]); k5b.append(t["query_b"])
    e1.record(stream)
    barrier()
    ms_total = max(e0.elapsed_time(e1), 0.0)
    wall_total = (time.perf_counter() - t0) * 1e3
    ms_total = max(ms_total, wall_total)
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    ms_step = ms_total / args.steps
    value = world * nq / (ms_step * 1e-3)
    e2e_value = world * nq / (ms_e2e / args.steps * 1e-3)

    # algorithmic bytes of one K5 launch (canonical counters, outside the timed region)
    cnt = ctx.count_d(nq, ro_d.data_ptr(), rd_d.data_ptr())
    k5_ms = float(np.mean(k5))
    peak, peak_src = measured_peak()
    achieved = cnt["bytes_algorithmic"] / (k5_ms * 1e-3) / 1e9
    fast_bytes = 48 * nq + 64 * cnt["fast_nodes"] + 16 * cnt["fast_disk_tests"] + 32 * cnt["hits"]
    traffic = None
    tp = os.path.join(ROOT, "profiles", "k5_traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(args.workload)
        except Exception:
            traffic = None

    if rank != 0:
        ctx.close()
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1 and not args.no_cpu:
        cpu = cpu_baseline_port(ctx, wl, ro_h, rd_h, float(n_paths), args.cpu_seconds)
    line = {
        "metric": "beam_queries_per_s", "value": value, "unit": "queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "photons": int(n_photons), "photon_paths": int(n_paths),
                   "rays_per_gpu": int(nq), "resolution": [W, H], "radius": radius, "grid": wl["grid"],
                   "max_leaf_photons": ctx_leaf(args), "l2": "inputs larger than L2 (photon SoA + tree >> 126 MB at 10M)",
                   "step": "K1 trace + [exchange] + K2-4 build + K6 tau + K5 query + K7 film",
                   "exchange": (args.exchange if world > 1 else None),
                   "mean_hits_per_ray": cnt["hits"] / nq, "mean_V": cnt["nodes_visited"] / nq,
                   "mean_T": cnt["disk_tests"] / nq, "mdisk_tests_per_s": cnt["disk_tests"] / (k5_ms * 1e-3) / 1e6,
                   "phase_ms": {"trace": float(np.mean(ktrace)), "build": float(np.mean(kbuild)),
                                "tau": float(np.mean(ktau)), "query": k5_ms, "query_k5a_traverse": float(np.mean(k5a)),
                                "query_k5b_leaf_pairs": float(np.mean(k5b)), "film": float(np.mean(kfilm))},
                   "fast_nodes_per_ray": cnt["fast_nodes"] / nq, "fast_disk_tests_per_ray": cnt["fast_disk_tests"] / nq,
                   "leaf_pairs_per_ray": cnt["leaf_pairs"] / nq, "pair_queue_overflow": cnt["pair_queue_overflow"]},
        "e2e": {"value": e2e_value, "unit": "queries/s", "h2d_bytes_per_step": int(nq * 28),
                "d2h_bytes_per_step": int(nq * 12), "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches_per_step(world, wl) * args.steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "kernel": "K5 = k_traverse<0,false> + k_leaf_fat<0,false,8> (+ k_leaf_pairs<0,false,8> for incoherent packets)",
                     "kernel_ms": k5_ms, "algorithmic_bytes": int(cnt["bytes_algorithmic"]),
                     "own_structure_bytes": int(fast_bytes), "own_structure_gbs": fast_bytes / (k5_ms * 1e-3) / 1e9,
                     "peak_source": peak_src},
        "clocks": sampler.summary(),
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def ctx_leaf(args):
    return args.leaf if args.leaf > 0 else 8


def launches_per_step(world, wl=None):
    """Kernels of libbre.so launched per step (profiles/r01_launches_*.csv lists them):
    trace: k_trace_flat, k_scan_reduce, k_scan_tiles, k_scan_apply, k_compact (per chunk of <= 4M paths) 5
    build: k_bounds_init, k_bounds, k_morton, 4 x (k_rs_hist, 3 scan launches, k_rs_scatter),
           k_gather, k_karras, k_refit_direct, k_refit_climb, k_emit_fast                             28
    query: k_tau_table (grid media), k_store_batch_args, then per ray batch k_traverse, k_leaf_fat,
           k_leaf_pairs (returns at once when every packet was coherent)                              5
    film:  k_film_add                                                                                 1
    Counted for one trace chunk and one ray batch (the 10 M-photon 1080p step); larger workloads launch
    more (one set per chunk / batch), so this is a lower bound there.  NCCL's own kernels are not counted."""
    tau = 1 if (wl is None or wl.get("grid")) else 0
    return 5 + 28 + 4 + tau + 1


def cpu_baseline_port(ctx, wl, ro_h, rd_h, n_paths, seconds):
    """The CPU oracle (restatement of the reference algorithm) on a bounded sample of the same
    rays against the same photons, all host cores."""
    orc = ge.load_orc()
    o = orc.Oracle()
    pos, pw, dr, rad = ctx.download_photons()
    return _cpu_run(o, pos, pw, dr, rad, ro_h, rd_h, n_paths, seconds, wl)


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


def run_reference(args):
    """CPU arm: the oracle port times the same workload on the host cores.  (oracle/_ref's
    reference-code harness is used when it has been built.)"""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    orc = ge.load_orc()
    wl = WORKLOADS[args.workload]
    o = orc.Oracle()
    set_medium(o, wl)
    lights = scene_lights(wl)
    vol = float(np.prod(BOX_HI - BOX_LO))
    radius = math.sqrt(wl["hits"] * vol / (math.pi * wl["photons"]))
    # photon pass on the CPU (single thread, like the reference's NumTasks=1 photon pass)
    t0 = time.perf_counter()
    ppp_pos, _, _, _ = o.trace(lights, BOX_LO, BOX_HI, 0, 200_000, MAX_DEPTH)
    n_paths = int(math.ceil(wl["photons"] / (ppp_pos.shape[0] / 200_000.0)))
    pos, pw, dr, _ = o.trace(lights, BOX_LO, BOX_HI, 0, n_paths, MAX_DEPTH)
    t_trace = time.perf_counter() - t0
    ro_h, rd_h = camera_rays(wl["w"], wl["h"])
    nq = ro_h.shape[0]
    rad = np.full(pos.shape[0], radius, np.float32)
    o.set_photons(pos, pw, dr, rad)
    t0 = time.perf_counter()
    o.build()
    t_build = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(1)
    per_step = max(1024, int(args.ref_rays))
    times = []
    for i in range(args.warmup + args.steps):
        idx = rng.integers(0, nq, per_step)
        t0 = time.perf_counter()
        o.query(ro_h[idx], rd_h[idx], float(n_paths), threads=cores)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    # a full step = photon pass + build + nq queries; scale the sampled query time to nq rays
    q_rate = per_step / float(np.mean(times))
    step_s = t_trace + t_build + nq / q_rate
    value = nq / step_s
    line = {"impl": "reference", "metric": "beam_queries_per_s", "value": value, "unit": "queries/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "photons": int(pos.shape[0]), "photon_paths": n_paths,
                       "rays_per_gpu": nq, "resolution": [wl["w"], wl["h"]], "radius": radius},
            "cpu_baseline": {"value": value, "unit": "queries/s", "cores": cores, "kind": "port",
                             "sample": f"photon pass {t_trace:.1f} s (1 thread) + tree build {t_build:.1f} s (1 thread) measured once; "
                                       f"query rate {q_rate:.0f}/s from {per_step} random rays per step on {cores} threads, "
                                       f"extrapolated to the {nq}-ray step"},
            "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="smoke10m_1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--leaf", type=int, default=0)
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-tile", action="store_true", help="K5 packet walk instead of the image-tile walk (A/B)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--ref-rays", type=int, default=65536)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
