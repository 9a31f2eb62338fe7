"""BASELINE.json's configs at their STATED sizes, with the library's options left on auto, against the oracle.

The small-size parity tests (test_gpu_parity.py) cover the arithmetic; these cover the launch shapes the quoted numbers
come from: which kernel the launcher picks by itself at 100 k / 320 k / 1 M instances (frame-coherent sort, register
allocation, prefetch depth) and at config 5's bench point.  Inputs are piccolo_b200/workloads.py's - the very arrays
bench.py runs.  Reference behaviour per instance: function/animation/skeleton.cpp:39-156; engine/shader/glsl/mesh.vert:72-111.
"""
import numpy as np
import pytest

from piccolo_b200 import assets as A
from piccolo_b200 import workloads as W
from test_gpu_parity import MAT_TOL, POS_TOL, assert_mats_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from piccolo_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()


def _download_ranges(b, ranges):
    return np.concatenate([b.download_palette(first=f, count=n) for f, n in ranges], 0)


def test_config4_one_million_mixed_clips_auto_sorted_vs_oracle(ctx, oracle, player):
    """1 M instances, 8 mixed clips, random phases (configs[3]).  At this size the launcher switches the frame-coherent
    visiting order on BY ITSELF (three order_* launches before the pose kernel) and uses the 255-register, two-bones-ahead
    form; 4 096 instances in 64 sub-ranges - the first and the last instance included - are held against the oracle."""
    sk, cl, am, mk = player
    clips = W.config4_clips(cl)
    skel = ctx.upload_skeleton(sk)
    ids = np.int32([ctx.upload_clip(skel, c, am) for c in clips])
    N = 1000000
    pick, phase = W.config4_inputs(0, N)
    b = ctx.batch(skel, N, 1)
    try:
        b.set(clip_ids=ids[pick], phase=phase)
        l0 = ctx.launch_count()
        b.pose()
        assert ctx.launch_count() - l0 == 4, "expected order_hist + order_scan + order_scatter + pose at 1 M instances on auto"
        ranges = W.sample_ranges(N, 64, 64)
        got = _download_ranges(b, ranges)
        sel = np.concatenate([np.arange(f, f + n) for f, n in ranges])
        assert sel[0] == 0 and sel[-1] == N - 1 and len(sel) == 4096
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(c, am) for c in clips], pick[sel], phase[sel])
        assert_mats_close(got, ref)
        # tick on the same launch shape: phases advance exactly like animation_component.cpp:19-21 (bit-exact)
        lens = np.float32([c.total_frame / 30.0 for c in clips])[pick]
        b.set(clip_len=lens)
        b.tick(1.0 / 60.0, 2)
        ph2 = np.concatenate([b.phases(first=f, count=n) for f, n in ranges[:8]], 0)
        s8 = sel[:len(ph2)]
        want = phase[s8] + np.float32(1.0 / 60.0) / lens[s8]
        want = want - np.floor(want)
        assert np.array_equal(ph2, want.astype(np.float32))
        got2 = _download_ranges(b, ranges[:8])
        ref2 = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(c, am) for c in clips], pick[s8], ph2)
        assert_mats_close(got2, ref2)
    finally:
        b.close()


def test_config4_one_shard_of_eight_auto_sorted_vs_oracle(ctx, oracle, player):
    """configs[3] as one of eight ranks sees it: instances [375 000, 500 000) of the 1 M crowd.  With one clip per instance the
    launcher turns the frame-coherent order on from 49 152 instances: 4 launches, 2 048 sampled instances against the oracle."""
    sk, cl, am, mk = player
    clips = W.config4_clips(cl)
    skel = ctx.upload_skeleton(sk)
    ids = np.int32([ctx.upload_clip(skel, c, am) for c in clips])
    first, N = 375000, 125000
    pick, phase = W.config4_inputs(first, N)
    b = ctx.batch(skel, N, 1)
    try:
        b.set(clip_ids=ids[pick], phase=phase)
        l0 = ctx.launch_count()
        b.pose()
        assert ctx.launch_count() - l0 == 4, "expected the three order_* launches + pose at 125 000 one-clip instances on auto"
        ranges = W.sample_ranges(N, 32, 64, seed=9)
        got = _download_ranges(b, ranges)
        sel = np.concatenate([np.arange(f, f + n) for f, n in ranges])
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(c, am) for c in clips], pick[sel], phase[sel])
        assert_mats_close(got, ref)
    finally:
        b.close()


def test_two_clip_blend_at_320k_auto_launch_shape_vs_oracle(ctx, oracle, player):
    """K = 2 beyond 300 000 instances: the launcher sorts on the (bucket of clip 0, bucket of clip 1) pair and takes the
    3-blocks-per-SM register allocation.  config 3's inputs, 4 096 sampled instances against the oracle (A4 blend extension:
    the oracle is the definition)."""
    sk, cl, am, mk = player
    clips = W.config3_clips(cl)
    skel = ctx.upload_skeleton(sk)
    ids = np.int32([ctx.upload_clip(skel, c, am) for c in clips])
    N = 320000
    phase, weight = W.config3_inputs(0, N)
    b = ctx.batch(skel, N, 2)
    try:
        b.set(clip_ids=np.tile(ids, (N, 1)), phase=phase, weight=weight)
        l0 = ctx.launch_count()
        b.pose()
        assert ctx.launch_count() - l0 == 4
        ranges = W.sample_ranges(N, 64, 64, seed=5)
        got = _download_ranges(b, ranges)
        sel = np.concatenate([np.arange(f, f + n) for f, n in ranges])
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(c, am) for c in clips], np.tile(np.int32([0, 1]), (len(sel), 1)),
                                phase[sel], weight=weight[sel])
        assert_mats_close(got, ref, tol=2e-5)
    finally:
        b.close()


def test_config3_full_size_256_instances_over_all_ring_chunks(ctx, oracle, player):
    """configs[2] exactly as bench.py runs it: 100 000 instances, two-clip blend, 16 380 vertices, a 25 024-slot ring filled
    by four launches (the last one 24 928 instances: 779 whole tiles).  64 instances of every chunk - 256 in all, first and
    last slot of each chunk included - are compared with the oracle: palette, positions, normals, tangents."""
    from piccolo_b200 import capi
    sk, cl, am, mk = player
    clips = W.config3_clips(cl)
    mesh = A.make_player_mesh(sk, V=16380, seed=42)
    skel = ctx.upload_skeleton(sk)
    ids = np.int32([ctx.upload_clip(skel, c, am) for c in clips])
    mh = ctx.upload_mesh(mesh, sk.J)
    N, ring = 100000, 25024
    phase, weight = W.config3_inputs(0, N)
    b = ctx.batch(skel, N, 2)
    sb = ctx.skinbuf(mh, mesh.V, ring, capi.PCL_SKIN_POS_NRM_TAN)
    try:
        b.set(clip_ids=np.tile(ids, (N, 1)), phase=phase, weight=weight)
        b.tick(0.0, capi.PCL_ADVANCE_ALL)  # the bench's call; dt = 0 keeps the phases
        sv, ocl = oracle.skeleton(sk), [oracle.clip(c, am) for c in clips]
        worst_p = worst_n = 0.0
        for ci, f in enumerate(range(0, N, ring)):
            n = min(ring, N - f)
            b.skin(mh, sb, first=f, count=n, slot_first=0)
            slots = [0] + sorted(int(x) + 1 for x in np.random.default_rng(20 + ci).choice(n - 2, 62, replace=False)) + [n - 1]
            inst = np.int64(slots) + f
            ref = oracle.pose_batch(sv, ocl, np.tile(np.int32([0, 1]), (64, 1)), phase[inst], weight=weight[inst])
            for k, slot in enumerate(slots):
                pal = b.download_palette(first=int(inst[k]), count=1)
                assert_mats_close(pal, ref[k:k + 1], tol=2e-5)
                pos, nt = sb.download(slot, 1)
                rp, rn, rt = oracle.lbs(mesh, oracle.palette_slots(ref[k]))
                ep = np.abs(pos[0] - rp).max() / max(1.0, np.abs(rp).max())
                en = max(np.abs(nt[0][:, :3] - rn).max(), np.abs(nt[0][:, 3:] - rt).max())
                worst_p, worst_n = max(worst_p, ep), max(worst_n, en)
                assert ep <= POS_TOL and en <= 1e-4, "chunk %d slot %d: dp %.3g dn %.3g" % (ci, slot, ep, en)
        print("config3 full size: 256 instances, worst dp %.2e dn %.2e" % (worst_p, worst_n))
    finally:
        sb.close(), b.close()


@pytest.mark.parametrize("coherent", [False, True, "chain"])
def test_config5_at_its_bench_point_n512(ctx, oracle, coherent):
    """configs[4] at the bench's N = 512 (256 joints, 49 998 vertices, POS+NRM+TAN), options on auto: 32 of the 512
    instances - first and last included - against the oracle, on the random-binding mesh, on the nearest-four mesh and on
    the limb-bound one."""
    from piccolo_b200 import capi
    sk, cl, am, mesh = W.config5_assets(coherent)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    mh = ctx.upload_mesh(mesh, sk.J)
    N = 512
    ph = W.config5_phases(N)
    b = ctx.batch(skel, N, 1)
    sb = ctx.skinbuf(mh, mesh.V, N, capi.PCL_SKIN_POS_NRM_TAN)
    try:
        b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
        b.pose()
        b.skin(mh, sb)
        sel = sorted({0, N - 1} | set(range(7, N, 17)))[:32]
        sel[-1] = N - 1
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((len(sel), 1), np.int32), ph[sel])
        for k, i in enumerate(sel):
            pal = b.download_palette(first=i, count=1)
            assert_mats_close(pal, ref[k:k + 1], tol=2e-5)
            pos, nt = sb.download(i, 1)
            rp, rn, rt = oracle.lbs(mesh, oracle.palette_slots(ref[k]))
            assert np.abs(pos[0] - rp).max() <= POS_TOL * max(1.0, np.abs(rp).max())
            assert np.abs(nt[0][:, :3] - rn).max() <= 1e-4 and np.abs(nt[0][:, 3:] - rt).max() <= 1e-4
    finally:
        sb.close(), b.close()


def test_config2_ten_thousand_instances_auto_vs_oracle(ctx, oracle, player):
    """configs[1]: 10 000 instances, single clip, phase_i = frac(i * 0.618...), every instance against the oracle."""
    sk, cl, am, mk = player
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    N = 10000
    ph = W.config2_inputs(0, N)
    b = ctx.batch(skel, N, 1)
    try:
        b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
        b.pose()
        got = b.download_palette()
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), ph)
        assert_mats_close(got, ref, tol=MAT_TOL)
    finally:
        b.close()
