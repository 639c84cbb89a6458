#!/usr/bin/env python
"""tools/check_photon_shard.py -- run under torchrun on R >= 2 GPUs of one box.

The photon-sharded render (bre_set_photon_sharding + bre_film_gather) against a single-rank render of the same
frame: every rank traces its slice of the photon paths and KEEPS it, builds the LBVH of that slice, answers every
ray from it, and the partial films are summed into rank 0's image.  Checks, per rank:
  * the rank's store equals its slice of the single-rank store bit for bit (slices in rank order = the whole store);
  * per-ray hit counts summed over the ranks equal the single-rank counts (the hit SETS partition by photon);
  * rank 0's image equals the single-rank image within 1e-4 mean / 1e-3 max (float summation order);
  * fed with partials whose sum is exact in float (pixel p's single-rank radiance on rank p % R, zero elsewhere),
    the gather (P2P reduce-scatter and ncclReduce) reproduces the single-rank image bit for bit.
Prints one line per rank, exits 1 on any mismatch.
"""
import ctypes
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bre = ge.load_bre()
    wl = dict(bench.WORKLOADS["smoke10m_1080p"], photons=400000)

    def make_ctx(no_p2p=False):
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            raw = ctypes.create_string_buffer(128)
            assert bre.load_library().bre_nccl_unique_id(raw) == 0
            buf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).cuda()
        dist.broadcast(buf, 0)
        if no_p2p:
            os.environ["BRE_NO_P2P"] = "1"
        try:
            c = bre.Bre(device=local, n_ranks=world, rank=rank, nccl_unique_id=bytes(buf.cpu().numpy().tobytes()))
        finally:
            os.environ.pop("BRE_NO_P2P", None)
        bench.setup_scene(c, wl)
        return c

    W, H = 203, 121            # ragged tiles, pixel count not a multiple of 4 or of the rank count
    radius, n_paths = 0.012, 150000
    ro, rd = bench.camera_rays(W, H)
    solo = bre.Bre(device=local)
    bench.setup_scene(solo, wl)
    solo.trace_photons(0, n_paths, bench.MAX_DEPTH, radius)
    solo.build()
    solo.set_ray_grid(W, H)
    s_rgb, s_offs, s_idx = solo.query(ro, rd, float(n_paths), want_hits=True)
    t = torch.from_numpy(s_rgb).cuda()
    dist.broadcast(t, 0)
    s_rgb0 = t.cpu().numpy()
    solo.film_create(W * H)
    solo.film_accumulate(s_rgb0)
    s_img = solo.film_resolve(0, 1.0)
    s_mean = solo.film_mean(0)
    s_store = solo.download_photons()
    s_cnt = np.diff(s_offs.astype(np.int64))

    ok = True
    imgs = []
    for no_p2p in (False, True):
        c = make_ctx(no_p2p)
        c.set_photon_sharding(True)
        n_local = c.trace_exchange_photons(n_paths, bench.MAX_DEPTH, radius)
        counts = torch.zeros(world, dtype=torch.int64, device="cuda")
        counts[rank] = int(n_local)
        dist.all_reduce(counts)
        counts = counts.cpu().numpy()
        first = int(counts[:rank].sum())
        mine = c.download_photons()
        store_ok = int(counts.sum()) == s_store[0].shape[0] and all(
            np.array_equal(a, b[first:first + int(n_local)]) for a, b in zip(mine, s_store))
        c.build()
        c.set_ray_grid(W, H)
        rgb, offs, idx = c.query(ro, rd, float(n_paths), want_hits=True)
        cnt = torch.from_numpy(np.diff(offs.astype(np.int64))).cuda()
        dist.all_reduce(cnt)
        hits_ok = np.array_equal(cnt.cpu().numpy(), s_cnt)
        c.film_create(W * H)
        c.film_accumulate(rgb)
        img = c.film_gather(0, 1.0, want=(rank == 0))
        img_ok = True
        why = ""
        if rank == 0:
            err = np.abs(img.astype(np.float64) - s_img)
            img_ok = err.mean() <= 1e-4 * np.abs(s_img).mean() and err.max() <= 1e-3 * np.abs(s_img).max()
            if not img_ok:
                why += f" render mean {err.mean() / np.abs(s_img).mean():.2e} max {err.max() / np.abs(s_img).max():.2e}"
        img2 = c.film_gather(0, 1.0, want=(rank == 0))          # other half of rank 0's double buffer
        if rank == 0 and not np.array_equal(img, img2):
            why += " second-gather"
            img_ok = False
        # exact partials: pixel p lives on rank p % R only -> the sum is exact, the image must match bit for bit
        sel = (np.arange(W * H) % world) == rank
        c.film_clear()
        c.film_accumulate(s_rgb0 * sel[:, None])
        img3 = c.film_gather(0, 1.0, want=(rank == 0))
        mean3 = c.film_gather(0, 1.0, want=(rank == 0), mean_only=True)
        if rank == 0:
            if not np.array_equal(img3, s_img):
                bad = np.nonzero((img3 != s_img).any(axis=1))[0]
                why += f" exact-gather {bad.size} px first {bad[:5].tolist()}"
                img_ok = False
            if not np.array_equal(mean3, s_mean):
                why += " exact-mean"
                img_ok = False
            imgs.append(img3)
        print(f"rank {rank} no_p2p={no_p2p}: store {store_ok} hits {hits_ok} image {img_ok} "
              f"(photons {int(n_local)} of {int(counts.sum())}){why}", flush=True)
        ok = ok and store_ok and hits_ok and img_ok
        dist.barrier()
        c.close()
    if rank == 0:
        same = np.array_equal(imgs[0], imgs[1])
        print(f"rank 0: P2P gather == NCCL gather {same}", flush=True)
        ok = ok and same
    solo.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
