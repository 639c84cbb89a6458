"""bre_film_resolve_static: the allocation-free variant of the film resolve (include/bre.h), through the C ABI.

Collected last on purpose (file name).  The static device scratch is shared by every
call in the process and keeps the last call's contents, so the test runs two calls in a row on different films and
checks each against the allocating entry points bre_film_resolve / bre_film_mean -- same kernel, so bit for bit, which
is inside the radiance tolerance the other tests use (1e-4 mean / 1e-3 max).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MEAN_TOL, MAX_TOL = 1e-4, 1e-3


def _close(a, b):
    err = np.abs(a.astype(np.float64) - b)
    assert err.mean() <= MEAN_TOL * max(np.abs(b).mean(), 1e-30)
    assert err.max() <= MAX_TOL * max(np.abs(b).max(), 1e-30)
    assert np.array_equal(a, b)          # the same kernel wrote both


def test_film_resolve_static_matches_film_resolve_two_calls_in_a_row(bre):
    rng = np.random.default_rng(7)
    ctx = bre.Bre(device=0)
    try:
        # first call: a 333 x 217 film (pixel count not a multiple of the block size), one iteration
        n1 = 333 * 217
        ctx.film_create(n1)
        ctx.film_accumulate((rng.random((n1, 3), dtype=np.float32) * 3 - 0.5).astype(np.float32))   # some negative: max(0, .)
        got1 = ctx.film_resolve_static(0, 1.7)
        _close(got1, ctx.film_resolve(0, 1.7))
        # second call right after it, on a SMALLER film with other contents and two iterations: nothing of the first
        # call's triples (still in the scratch) may show
        n2 = 97 * 61
        ctx.film_create(n2)
        for _ in range(2):
            ctx.film_accumulate(rng.random((n2, 3), dtype=np.float32))
        got2 = ctx.film_resolve_static(1, 0.5)
        assert got2.shape == (n2, 3)
        _close(got2, ctx.film_resolve(1, 0.5))
        # mean_only = 1 is bre_film_mean
        _close(ctx.film_resolve_static(1, 1.0, mean_only=True), ctx.film_mean(1))
        # and the first film again on a second context of the same device (same scratch, address looked up once)
        other = bre.Bre(device=0)
        try:
            other.film_create(n2)
            other.film_accumulate(np.full((n2, 3), 0.25, np.float32))
            _close(other.film_resolve_static(0, 2.0), other.film_resolve(0, 2.0))
        finally:
            other.close()
        _close(ctx.film_resolve_static(1, 0.5), got2)
    finally:
        ctx.close()


def test_film_resolve_static_rejects_films_beyond_the_static_scratch(bre):
    ctx = bre.Bre(device=0)
    try:
        ctx.film_create((1 << 21) + 1)
        with pytest.raises(bre.BreError) as e:
            ctx.film_resolve_static(0, 1.0)
        assert "BRE_STATIC_RESOLVE_MAX_PIXELS" in str(e.value)
        ctx.film_create(1 << 21)                       # exactly the maximum is accepted
        ctx.film_accumulate(np.full((1 << 21, 3), 0.5, np.float32))
        _close(ctx.film_resolve_static(0, 1.0), ctx.film_resolve(0, 1.0))
    finally:
        ctx.close()
