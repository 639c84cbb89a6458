"""CPU tests of the drop-in boundary: libbre.so loads, exports every symbol include/bre.h declares,
refuses to run without a CUDA device (no CPU fallback), and its pure host-side sharding helpers work.
No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "bre.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bre_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_list_agree(bre):
    assert header_symbols() == sorted(bre.SYMBOLS)


def test_library_exports_every_declared_symbol(bre):
    lib = bre.load_library()
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    assert lib.bre_abi_version() == 3


def test_no_cpu_fallback(bre):
    """Without a CUDA device bre_create must fail with BRE_ERR_NO_DEVICE; with one it must succeed."""
    lib = bre.load_library()
    ctx = C.c_void_p()
    cfg = bre.Config(0, 1, 0, None, 0, 0)
    s = lib.bre_create(C.byref(ctx), C.byref(cfg))
    import torch
    if torch.cuda.is_available():
        assert s == 0
        lib.bre_destroy(ctx)
    else:
        assert s == -2 and not ctx.value
        try:
            bre.Bre(device=0)
        except bre.BreError as e:
            assert e.code == -2
        else:
            raise AssertionError("Bre() must raise without a CUDA device")


def test_null_context_is_rejected(bre):
    lib = bre.load_library()
    assert lib.bre_build(None) == -1
    assert lib.bre_query(None, 0, None, None, C.c_float(1), None, None) == -1
    assert lib.bre_destroy(None) in (0, -1)


def test_shard_range_partitions_exactly(bre):
    lib = bre.load_library()
    for n in (0, 1, 7, 4096, 3_907_372, (1 << 33) + 5):
        for ranks in (1, 2, 3, 8):
            covered = 0
            prev_end = 0
            sizes = []
            for r in range(ranks):
                first, count = bre.shard_range(n, ranks, r, lib)
                assert first == prev_end          # contiguous, rank-major
                prev_end = first + count
                covered += count
                sizes.append(count)
            assert covered == n and prev_end == n
            assert max(sizes) - min(sizes) <= 1   # balanced


def test_tile_owner_round_robin(bre):
    lib = bre.load_library()
    tiles_x, tiles_y = 120, 68                    # 1920x1080 in 16x16 tiles (photonbeam.cpp:345-347)
    for ranks in (1, 2, 4, 8):
        own = np.array([[bre.tile_owner(tx, ty, tiles_x, ranks, lib) for tx in range(tiles_x)]
                        for ty in range(tiles_y)])
        assert own.min() == 0 and own.max() == ranks - 1
        counts = np.bincount(own.ravel(), minlength=ranks)
        assert counts.max() - counts.min() <= 1
