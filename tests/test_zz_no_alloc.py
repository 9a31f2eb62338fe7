"""pcl_batch_download_palette_static -- the read-back entry point that allocates no device memory.

It exists for callers that cannot call cudaMalloc: code inside a frame loop (an allocation would stall every
stream on the GPU for an unbounded time) and hosts whose own allocator already owns the whole device-memory
budget.  The call stages through a fixed 8 MiB array inside the library instead of a cudaMalloc'ed buffer, so
its result must be the same matrices pcl_batch_download_palette returns, call after call (the array is shared
and keeps the previous call's contents), and ranges beyond the array must be refused rather than truncated.

This file sorts last on purpose: a run that stops at the first failure has then executed every other test.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_static_download_matches_regular_download_twice_in_a_row(player):
    from piccolo_b200 import capi
    sk, cl, am, mk = player
    ctx = capi.Context(0)
    try:
        skel = ctx.upload_skeleton(sk)
        clip = ctx.upload_clip(skel, cl, am)
        N = 3000
        ph = np.random.default_rng(2024).uniform(0, 1, N).astype(np.float32)
        b = ctx.batch(skel, N, 1)
        b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
        b.pose()
        want = b.download_palette()  # existing entry point (cudaMalloc'ed staging buffer)
        limit = capi.PCL_STATIC_EXPORT_MAX_BYTES // (sk.J * 64)  # 2259 instances of the player skeleton
        assert 0 < limit < N

        # two calls in a row over different, overlapping ranges: the second must not see the first one's leftovers
        first_call = b.download_palette_static(0, limit)
        second_call = b.download_palette_static(N - 1500, 1500)
        for got, ref in ((first_call, want[:limit]), (second_call, want[N - 1500:])):
            scale = np.maximum(1.0, np.abs(ref).max(axis=2, keepdims=True))
            assert (np.abs(got - ref) / scale).max() <= 1e-5  # the tolerance the palette tests use
            assert np.array_equal(got, ref)  # same kernel, same palette: in fact identical

        # a new pose, then the same range again: fresh matrices, not the stale contents of the array
        b.set(phase=np.modf(ph.astype(np.float64) + 0.25)[0].astype(np.float32))
        b.pose()
        again = b.download_palette_static(N - 1500, 1500)
        assert np.array_equal(again, b.download_palette(N - 1500, 1500))
        assert not np.array_equal(again, second_call)

        # beyond the fixed array: refused with a clear message, nothing written
        with pytest.raises(capi.PclError) as e:
            b.download_palette_static(0, limit + 1)
        assert e.value.code == -3 and "PCL_STATIC_EXPORT_MAX_BYTES" in str(e.value)
        b.close()
    finally:
        ctx.close()
