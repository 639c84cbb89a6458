#!/usr/bin/env python
"""tools/check_exchange.py -- run under torchrun on R >= 2 GPUs of one box.

Checks the fused trace + push exchange (bre_trace_exchange_photons, CUDA-IPC writes over NVLink)
against (a) the NCCL all-gather path and (b) a single-rank trace of the same path indices: the
photon store must be BIT-IDENTICAL on every rank in all three (rank-major order = path-index
order), over several rounds with growing photon counts (forces a collective re-export of the
stores) and after a non-collective store growth in between.  Prints one line per rank, exits 1
on any mismatch.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tools/check_exchange.py
"""
import ctypes
import hashlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def digest(ctx):
    h = hashlib.sha256()
    for a in ctx.download_photons():
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def scene(ctx):
    ctx.set_medium_homogeneous([0.05, 0.06, 0.07], [0.4, 0.35, 0.3], 0.3)
    ctx.set_lights([dict(kind=0, pos=(0.0, 0.9, 0.0), I=(5.0, 4.0, 3.0))])
    ctx.set_bounds([-4, -1, -4], [4, 2, 4])


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bre = ge.load_bre()

    def make_ctx():
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            raw = ctypes.create_string_buffer(128)
            assert bre.load_library().bre_nccl_unique_id(raw) == 0
            buf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).cuda()
        dist.broadcast(buf, 0)
        c = bre.Bre(device=local, n_ranks=world, rank=rank, nccl_unique_id=bytes(buf.cpu().numpy().tobytes()))
        scene(c)
        return c

    p2p, nccl = make_ctx(), make_ctx()
    solo = bre.Bre(device=local)
    scene(solo)
    ok = True
    rounds = [20000, 5000, 150000, 150001, 400000]       # growing counts force a re-export
    for i, n_paths in enumerate(rounds):
        n1 = p2p.trace_exchange_photons(n_paths, 5, 0.1)
        first, count = bre.shard_range(n_paths, world, rank, p2p.lib)
        nccl.clear_photons()
        nccl.trace_photons(first, count, 5, 0.1)
        nccl.allgather_photons()
        solo.clear_photons()
        n3 = solo.trace_photons(0, n_paths, 5, 0.1)
        d1, d2, d3 = digest(p2p), digest(nccl), digest(solo)
        same = d1 == d2 == d3 and n1 == nccl.photon_count() == n3
        ok &= same
        print(f"rank {rank} round {i}: {n1} photons p2p==nccl==solo: {same}", flush=True)
        if i == 1:   # non-collective growth of an exported store on one rank only
            if rank == 0:
                z = np.zeros((300000, 3), np.float32)
                p2p.upload_photons(z, z, z, radius=0.1, append=True)
    # the exchanged store feeds the normal pipeline
    p2p.build()
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    p2p.close(); nccl.close(); solo.close()
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


if __name__ == "__main__":
    main()
