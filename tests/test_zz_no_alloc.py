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

        # the point of the entry point: once its one-time address look-up is behind it, no call - served or refused - takes
        # or returns a single byte of device memory (a version that quietly fell back to a cudaMalloc'ed buffer would)
        import torch
        torch.cuda.init()
        ctx.sync()
        free0 = torch.cuda.mem_get_info(0)[0]
        for k in range(6):
            part = b.download_palette_static(100 * k, limit - 200 * k)
            assert np.array_equal(part, b.download_palette_static(100 * k, limit - 200 * k))
            with pytest.raises(capi.PclError):
                b.download_palette_static(0, limit + 1 + k)
            assert torch.cuda.mem_get_info(0)[0] == free0, "device memory changed hands in call %d" % k

        # a second context on the same GPU shares the one array (and its address): same matrices through it
        ctx2 = capi.Context(0)
        try:
            skel2 = ctx2.upload_skeleton(sk)
            clip2 = ctx2.upload_clip(skel2, cl, am)
            b2 = ctx2.batch(skel2, 500, 1)
            b2.set(clip_ids=np.full((500, 1), clip2, np.int32), phase=ph[:500])
            b2.pose()
            assert np.array_equal(b2.download_palette_static(0, 500), b2.download_palette(0, 500))
            b2.close()
        finally:
            ctx2.close()
        b.close()
    finally:
        ctx.close()
