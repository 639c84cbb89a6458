#!/usr/bin/env python
"""tools/check_tiles.py -- run under torchrun on R >= 2 GPUs of one box.

The tile-sharded camera pass (bre_set_tile_sharding + bre_film_gather) against a single-rank render of the same
frame: every rank traces its slice of the photon paths and receives everyone's (fused push), builds the LBVH,
answers ONLY the 16x16 tiles bre_tile_owner gives it, and rank 0 gathers the film.  Checks, per rank:
  * the photon store equals the single-rank store bit for bit;
  * per-ray hit-index sets: equal to the single-rank sets on owned tiles, empty elsewhere;
  * rank 0's gathered image equals the single-rank image within 1e-4 mean / 1e-3 max (float atomics);
  * fed with the single-rank radiance, the gather (P2P push and NCCL reduce) reproduces the single-rank image bit for bit.
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

    W, H = 200, 120            # 13 x 8 tiles, ragged right column and bottom row
    radius, n_paths = 0.012, 150000
    ro, rd = bench.camera_rays(W, H)
    solo = bre.Bre(device=local)
    bench.setup_scene(solo, wl)
    solo.trace_photons(0, n_paths, bench.MAX_DEPTH, radius)
    solo.build()
    solo.set_ray_grid(W, H)
    s_rgb, s_offs, s_idx = solo.query(ro, rd, float(n_paths), want_hits=True)
    # radiance carries the order of its float atomics: every rank uses RANK 0's single-rank radiance as the input
    # of the exact-gather check below
    t = torch.from_numpy(s_rgb).cuda()
    dist.broadcast(t, 0)
    s_rgb0 = t.cpu().numpy()
    solo.film_create(W * H)
    solo.film_accumulate(s_rgb0)
    s_img = solo.film_resolve(0, 1.0)
    s_store = solo.download_photons()

    ok = True
    imgs = []
    for no_p2p in (False, True):
        c = make_ctx(no_p2p)
        c.trace_exchange_photons(n_paths, bench.MAX_DEPTH, radius)
        same_store = all(np.array_equal(a, b) for a, b in zip(c.download_photons(), s_store))
        c.build()
        c.set_ray_grid(W, H)
        c.set_tile_sharding(True)
        rgb, offs, idx = c.query(ro, rd, float(n_paths), want_hits=True)
        tx = (np.arange(W * H) % W) // 16
        ty = (np.arange(W * H) // W) // 16
        mine = ((ty * ((W + 15) // 16) + tx) % world) == rank
        cnt, s_cnt = np.diff(offs.astype(np.int64)), np.diff(s_offs.astype(np.int64))
        hits_ok = np.array_equal(cnt[mine], s_cnt[mine]) and not cnt[~mine].any()
        if hits_ok:      # same index lists on the owned rays
            keep = np.repeat(mine, s_cnt)
            hits_ok = np.array_equal(idx, s_idx[keep])
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
            imgs.append(img)
        # a second gather (other half of rank 0's double buffer) gives the same image
        img2 = c.film_gather(0, 1.0, want=(rank == 0))
        if rank == 0:
            if not np.array_equal(img, img2):
                why += " second-gather"
            img_ok = img_ok and np.array_equal(img, img2)
        # the gather itself is exact: fed with the single-rank radiance of its own tiles, every rank's film
        # assembles to the single-rank image bit for bit
        c.film_clear()
        c.film_accumulate(s_rgb0 * mine[:, None])
        img3 = c.film_gather(0, 1.0, want=(rank == 0))
        if rank == 0:
            if not np.array_equal(img3, s_img):
                bad = np.nonzero((img3 != s_img).any(axis=1))[0]
                why += f" exact-gather {bad.size} px first {bad[:5].tolist()} {img3[bad[:2]].tolist()} vs {s_img[bad[:2]].tolist()}"
            img_ok = img_ok and np.array_equal(img3, s_img)
            imgs[-1] = img3
        print(f"rank {rank} no_p2p={no_p2p}: store {same_store} hits {hits_ok} image {img_ok} "
              f"(owned rays {int(mine.sum())}, hits {int(cnt.sum())}){why}", flush=True)
        ok = ok and same_store and hits_ok and img_ok
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
