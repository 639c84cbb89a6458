"""world_size-2 test of the multi-GPU partitioning on CPU (gloo): photon paths are split by index
with bre_shard_range, gathered rank-major, and the 16x16 image tiles are dealt with bre_tile_owner.
The compute on each rank is done by the CPU oracle (this is a test of the HOST-SIDE sharding logic,
the GPU path does the same with NCCL): the gathered photon set and the assembled image must be
bit-identical to the single-process result."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_PATHS, DEPTH, W, H, TILE = 1500, 5, 32, 32, 16
SA, SS, G = [0.05] * 3, [0.4] * 3, 0.3
LIGHTS = [dict(kind=0, pos=(0.5, 0.5, 0.5), I=(8, 8, 8))]
LO, HI = np.zeros(3, np.float32), np.ones(3, np.float32)
RADIUS = 0.08


def _setup():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    return ge.load_bre(), ge.load_orc()


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bre, orc = _setup()
    lib = bre.load_library()
    o = orc.Oracle()
    o.set_medium_homogeneous(SA, SS, G)
    # 1. photon paths split by index
    first, count = bre.shard_range(N_PATHS, world, rank, lib)
    pos, pw, dr, _ = o.trace(LIGHTS, LO, HI, first, count, DEPTH)
    rec = torch.from_numpy(np.concatenate([pos, pw, dr], axis=1).copy())
    # 2. all-gather: counts, then padded slabs, compacted rank-major (what bre_allgather_photons does)
    cnt = torch.tensor([rec.shape[0]], dtype=torch.int64)
    cnts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(cnts, cnt)
    mx = int(max(c.item() for c in cnts))
    pad = torch.zeros((mx, 9), dtype=torch.float32)
    pad[:rec.shape[0]] = rec
    slabs = [torch.zeros((mx, 9), dtype=torch.float32) for _ in range(world)]
    dist.all_gather(slabs, pad)
    allrec = torch.cat([slabs[r][:int(cnts[r].item())] for r in range(world)]).numpy()
    # 3. every rank builds the same tree and renders the tiles it owns
    o.set_photons(allrec[:, 0:3], allrec[:, 3:6], allrec[:, 6:9], RADIUS)
    o.build()
    ro, rd = o.synth_rays_pinhole(W, H)
    tiles_x = W // TILE
    px = np.arange(W * H)
    owner = np.array([bre.tile_owner((p % W) // TILE, (p // W) // TILE, tiles_x, world, lib) for p in px])
    mine = px[owner == rank]
    rgb, _ = o.query(ro[mine], rd[mine], float(N_PATHS))
    img = torch.zeros((W * H, 3), dtype=torch.float32)
    img[torch.from_numpy(mine)] = torch.from_numpy(rgb)
    # 4. disjoint tiles: a SUM reduce to rank 0 is an exact gather
    dist.reduce(img, 0, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.savez(out_path, photons=allrec, image=img.numpy(), owner=owner)
    dist.destroy_process_group()


def test_two_rank_sharding_equals_single_process(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    bre, orc = _setup()
    o = orc.Oracle()
    o.set_medium_homogeneous(SA, SS, G)
    pos, pw, dr, _ = o.trace(LIGHTS, LO, HI, 0, N_PATHS, DEPTH)
    ref = np.concatenate([pos, pw, dr], axis=1)
    assert got["photons"].shape == ref.shape
    assert np.array_equal(got["photons"].view(np.uint32), ref.view(np.uint32))
    o.set_photons(pos, pw, dr, RADIUS)
    o.build()
    ro, rd = o.synth_rays_pinhole(W, H)
    rgb, _ = o.query(ro, rd, float(N_PATHS))
    assert np.array_equal(got["image"].view(np.uint32), rgb.view(np.uint32))
    assert set(got["owner"].tolist()) == {0, 1}
    assert rgb.max() > 0


def _worker_photon_shard(rank, world, port, out_path):
    """Photon sharding (bre_set_photon_sharding): a rank keeps the paths it traced, builds its own tree, answers EVERY
    pixel from it; the partial films are summed onto rank 0 (the GPU path: reduce-scatter over CUDA-IPC / ncclReduce)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bre, orc = _setup()
    lib = bre.load_library()
    o = orc.Oracle()
    o.set_medium_homogeneous(SA, SS, G)
    first, count = bre.shard_range(N_PATHS, world, rank, lib)
    pos, pw, dr, _ = o.trace(LIGHTS, LO, HI, first, count, DEPTH)
    o.set_photons(pos, pw, dr, RADIUS)
    o.build()
    ro, rd = o.synth_rays_pinhole(W, H)
    rgb, offs, idx, _ = o.query(ro, rd, float(N_PATHS), want_hits=True)      # normalised by ALL emitted paths
    film = torch.from_numpy(rgb.astype(np.float32).copy())
    hits = torch.from_numpy(np.diff(offs.astype(np.int64)))
    n_local = torch.tensor([pos.shape[0]], dtype=torch.int64)
    dist.reduce(film, 0, op=dist.ReduceOp.SUM)
    dist.reduce(hits, 0, op=dist.ReduceOp.SUM)
    dist.reduce(n_local, 0, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.savez(out_path, image=film.numpy(), hits=hits.numpy(), n=n_local.numpy())
    dist.destroy_process_group()


def test_two_rank_photon_sharding_sums_to_the_single_process_image(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "summed.npz")
    mp.spawn(_worker_photon_shard, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    bre, orc = _setup()
    o = orc.Oracle()
    o.set_medium_homogeneous(SA, SS, G)
    pos, pw, dr, _ = o.trace(LIGHTS, LO, HI, 0, N_PATHS, DEPTH)
    assert int(got["n"][0]) == pos.shape[0]                       # the slices partition the photon set
    o.set_photons(pos, pw, dr, RADIUS)
    o.build()
    ro, rd = o.synth_rays_pinhole(W, H)
    rgb, offs, idx, _ = o.query(ro, rd, float(N_PATHS), want_hits=True)
    assert np.array_equal(got["hits"], np.diff(offs.astype(np.int64)))   # hit sets partition by photon
    err = np.abs(got["image"].astype(np.float64) - rgb)
    assert err.mean() <= 1e-4 * np.abs(rgb).mean() and err.max() <= 1e-3 * np.abs(rgb).max()   # float summation order
    assert rgb.max() > 0
