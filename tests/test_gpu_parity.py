"""Parity of the CUDA path (through the C ABI) against the oracle and the reference-generated goldens.

Tolerances (SURVEY.md section 0.7 / BASELINE.md section 3.5):
  joint matrices     max|d| <= 1e-5 * max(1, max|M_ref|)   per joint matrix
  skinned positions  |dp|   <= 1e-4 * max(1, bounding radius of the reference positions)
  normals/tangents   |dn|   <= 1e-4 (unit vectors)
  indices / ordering / phases: bit-exact
"""
import os

import numpy as np
import pytest

from piccolo_b200 import assets as A
from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu

MAT_TOL = 1e-5
POS_TOL = 1e-4


@pytest.fixture(scope="module")
def ctx():
    from piccolo_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def player_dev(ctx, player):
    sk, cl, am, mk = player
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    return skel, clip


def capi_mod():
    from piccolo_b200 import capi
    return capi


def assert_mats_close(got, ref, tol=MAT_TOL):
    got, ref = got.reshape(-1, 16), ref.reshape(-1, 16)
    scale = np.maximum(1.0, np.abs(ref).max(axis=1))
    err = np.abs(got.astype(np.float64) - ref.astype(np.float64)).max(axis=1) / scale
    assert np.all(np.isfinite(got)), "non-finite palette entries"
    assert err.max() <= tol, "joint-matrix error %.3g > %.1g at flat joint %d" % (err.max(), tol, int(err.argmax()))
    return float(err.max())


def gpu_pose(ctx, skel, clip_ids, phases, weight=None, mask_ids=None, derived=False, tick=None):
    clip_ids = np.asarray(clip_ids, np.int32)
    clip_ids = clip_ids.reshape(len(clip_ids), -1)
    N, K = clip_ids.shape
    b = ctx.batch(skel, N, K)
    try:
        b.set(clip_ids=clip_ids, phase=np.asarray(phases, np.float32).reshape(N, K), weight=weight, mask_ids=mask_ids)
        if derived:
            b.keep_derived(True)
        b.pose()
        mats = b.download_palette()
        return (mats, b.download_derived()) if derived else mats
    finally:
        b.close()


# ------------------------------------------------------------------------------------------------
# pose: goldens produced by the reference's own Skeleton code
# ------------------------------------------------------------------------------------------------
def test_pose_reference_goldens(ctx, player_dev):
    skel, clip = player_dev
    g = np.load(os.path.join(GOLDEN, "pose_golden.npz"))
    ph = g["phases"]
    mats, der = gpu_pose(ctx, skel, np.full((len(ph), 1), clip), ph, derived=True)
    assert_mats_close(mats, g["mats"])
    assert np.array_equal(mats[..., 12:], np.broadcast_to(np.float32([0, 0, 0, 1]), mats[..., 12:].shape))
    assert np.array_equal(ctx.skeleton_indices(skel), g["index"])  # bit-exact, bones_map order
    d = np.abs(der[g["derived_phase_idx"]] - g["derived"]).max()
    assert d <= 1e-5, "derived TRS error %.3g" % d
    # the root is exactly the identity in the reference
    assert np.array_equal(mats[:, 0], np.broadcast_to(np.eye(4, dtype=np.float32).reshape(16), (len(ph), 16)))


def test_pose_clamp_carry_over(ctx, player):
    g = np.load(os.path.join(GOLDEN, "clamp_golden.npz"))
    sk, cl, am, mk = A.load_player()
    for ch, track, keep in g["cut"]:
        tr = (cl.pos_keys, cl.rot_keys, cl.scl_keys)[track]
        tr[ch] = tr[ch][:keep].copy()
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    mats = gpu_pose(ctx, skel, np.full((len(g["phases"]), 1), clip), g["phases"])
    assert_mats_close(mats, g["mats"])


def test_tick_sequence_matches_animation_component(ctx, player_dev):
    """200 x AnimationComponent::tick: phases bit-exact, palettes within tolerance."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    g = np.load(os.path.join(GOLDEN, "tick_golden.npz"))
    b = ctx.batch(skel, 3, 1)
    b.set(clip_ids=np.full((3, 1), clip, np.int32), phase=np.zeros((3, 1), np.float32),
          clip_len=np.full((3, 1), g["clip_len"], np.float32))
    want = dict(zip(g["mats_tick_idx"].tolist(), g["mats"]))
    for t in range(len(g["phases"])):
        if t % 2:
            b.tick(float(g["dt"]), capi.PCL_ADVANCE_FIRST_ONLY)  # fused advance + pose
        else:
            b.advance(float(g["dt"]), capi.PCL_ADVANCE_FIRST_ONLY)  # separate kernels
            b.pose()
        ph = b.phases()
        assert np.all(ph == g["phases"][t]), "phase diverged at tick %d" % t
        if t in want:
            assert_mats_close(b.download_palette()[1], want[t])
    b.close()


# ------------------------------------------------------------------------------------------------
# pose: against the oracle on seeded inputs
# ------------------------------------------------------------------------------------------------
def test_pose_config2_10k_single_clip(ctx, player_dev, oracle, player):
    skel, clip = player_dev
    sk, cl, am, mk = player
    N = 10000
    ph = np.modf(np.arange(N, dtype=np.float64) * 0.6180339887)[0].astype(np.float32)
    mats = gpu_pose(ctx, skel, np.full((N, 1), clip), ph)
    ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), ph)
    assert_mats_close(mats, ref)


@pytest.mark.parametrize("N", [1, 31, 33, 1000, 4097])
def test_pose_ragged_batch_sizes(ctx, player_dev, oracle, player, N):
    skel, clip = player_dev
    sk, cl, am, mk = player
    ph = np.random.default_rng(N).uniform(0, 1, N).astype(np.float32)
    for wpb in (1, 2, 4, 8):
        ctx.set_option("pose_warps_per_block", wpb)
        mats = gpu_pose(ctx, skel, np.full((N, 1), clip), ph)
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), ph)
        assert_mats_close(mats, ref)
    ctx.set_option("pose_warps_per_block", 0)


def test_channel_ordered_pose_kernel_stays_correct(ctx, player_dev, oracle, player):
    """The generic (channel-ordered) kernel is what runs for clips with several channels per bone; force it on the
    bundled clip and on a blend so that both pose kernels are held to the same goldens."""
    skel, clip = player_dev
    sk, cl, am, mk = player
    g = np.load(os.path.join(GOLDEN, "pose_golden.npz"))
    ctx.set_option("pose_generic", 1)
    try:
        mats = gpu_pose(ctx, skel, np.full((len(g["phases"]), 1), clip), g["phases"])
        assert_mats_close(mats, g["mats"])
        other = ctx.upload_clip(skel, A.make_variant_clip(cl, 29, seed=5), am)
        rng = np.random.default_rng(3)
        ph, w = rng.uniform(0, 1, (500, 2)).astype(np.float32), rng.uniform(0.1, 1, (500, 2)).astype(np.float32)
        two = gpu_pose(ctx, skel, np.tile(np.int32([clip, other]), (500, 1)), ph, weight=w)
    finally:
        ctx.set_option("pose_generic", 0)
    same = gpu_pose(ctx, skel, np.tile(np.int32([clip, other]), (500, 1)), ph, weight=w)
    assert np.abs(two - same).max() <= 2e-6  # both kernels, same arithmetic: differ only by instruction scheduling/FMA


def _variants(cl, frames=(17, 23, 29, 31, 37, 41, 47)):
    return [cl] + [A.make_variant_clip(cl, f, seed=100 + i) for i, f in enumerate(frames)]


def test_pose_config4_mixed_clips_random_phases(ctx, oracle, player):
    sk, cl, am, mk = player
    clips = _variants(cl)
    skel = ctx.upload_skeleton(sk)
    ids = [ctx.upload_clip(skel, c, am) for c in clips]
    N = 20000
    rng = np.random.default_rng(7)
    pick = rng.integers(0, len(clips), N)
    ph = rng.uniform(0, 1, N).astype(np.float32)
    mats = gpu_pose(ctx, skel, np.asarray(ids, np.int32)[pick].reshape(N, 1), ph)
    sv = oracle.skeleton(sk)
    ref = oracle.pose_batch(sv, [oracle.clip(c, am) for c in clips], pick.reshape(N, 1).astype(np.int32), ph)
    assert_mats_close(mats, ref)


@pytest.mark.parametrize("K", [2, 3, 4])
def test_pose_blend_extension_vs_oracle(ctx, oracle, player, K):
    """Multi-clip blend (no executed reference counterpart; oracle = documented semantics A4)."""
    sk, cl, am, mk = player
    clips = _variants(cl, frames=(17, 29, 41))
    skel = ctx.upload_skeleton(sk)
    ids = np.asarray([ctx.upload_clip(skel, c, am) for c in clips], np.int32)
    rng = np.random.default_rng(K)
    masks = np.ones((3, sk.J), np.int32)
    masks[1, rng.choice(sk.J, 20, replace=False)] = 0
    masks[2, :] = 0
    masks[2, :5] = 1
    mids = [ctx.upload_mask(skel, m) for m in masks]
    N = 3000
    pick = np.stack([rng.permutation(len(clips))[:K] for _ in range(N)]).astype(np.int32)
    ph = rng.uniform(0, 1, (N, K)).astype(np.float32)
    w = rng.uniform(0.05, 1, (N, K)).astype(np.float32)
    w[::7, 0] = 0  # zero-weight slots
    mpick = rng.integers(0, 3, (N, K)).astype(np.int32)
    mpick[::5] = 0
    mats = gpu_pose(ctx, skel, ids[pick], ph, weight=w, mask_ids=np.asarray(mids, np.int32)[mpick])
    sv = oracle.skeleton(sk)
    ref = oracle.pose_batch(sv, [oracle.clip(c, am) for c in clips], pick, ph, weight=w, masks=masks, mask_ids=mpick)
    assert_mats_close(mats, ref, tol=2e-5)  # two renormalisations more than the single-clip path


def test_pose_multi_channel_per_bone_and_skipped_channels(ctx, oracle, player):
    """Two channels driving one bone apply in channel order; convert[n] = -1 skips a channel BEFORE the clamp."""
    sk, cl, am, mk = A.load_player()
    conv = am.convert.copy()
    conv[10] = conv[4]  # bone of channel 4 now also driven by channel 10; channel 10's old bone keeps the bind pose
    conv[20] = -1
    cl.rot_keys[20] = cl.rot_keys[20][:3].copy()  # would clamp later channels if it were not skipped
    cl.scl_keys[30] = cl.scl_keys[30][:11].copy()
    am2 = A.AnimSkelMap(conv)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am2)
    ph = np.linspace(0, 1, 257, dtype=np.float32)
    mats = gpu_pose(ctx, skel, np.full((257, 1), clip), ph)
    ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am2)], np.zeros((257, 1), np.int32), ph)
    assert_mats_close(mats, ref)


def test_pose_config5_synthetic_256_joints(ctx, oracle):
    sk = A.make_synthetic_skeleton(256)
    cl, am = A.make_synthetic_clip(sk, 32)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    N = 2000
    ph = np.random.default_rng(5).uniform(0, 1, N).astype(np.float32)
    mats, der = gpu_pose(ctx, skel, np.full((N, 1), clip), ph, derived=True)
    ref, rder = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), ph, derived=True)
    assert_mats_close(mats, ref, tol=2e-5)  # depth-8 chains of random rotations, entries O(1)
    assert np.abs(der - rder).max() <= 2e-5


def test_pose_wide_and_deep_hierarchies(ctx, oracle):
    """Star (every bone a child of the root: many parked parents avoided) and chain (parent always in registers)."""
    for parent_of in (lambda i: 0, lambda i: i - 1, lambda i: max(0, i - 3)):
        sk = A.make_synthetic_skeleton(64)
        sk.parent_index = np.array([A.INT_MAX] + [parent_of(i) for i in range(1, 64)], np.int32)
        cl, am = A.make_synthetic_clip(sk, 9, seed=3)
        skel = ctx.upload_skeleton(sk)
        clip = ctx.upload_clip(skel, cl, am)
        ph = np.linspace(0, 1, 100, dtype=np.float32)
        mats = gpu_pose(ctx, skel, np.full((100, 1), clip), ph)
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((100, 1), np.int32), ph)
        assert_mats_close(mats, ref, tol=5e-5)  # 63-deep chain of random rotations
        assert np.array_equal(ctx.skeleton_parents(skel), np.array([-1] + [parent_of(i) for i in range(1, 64)]))


@pytest.mark.parametrize("K", [1, 2, 3])
@pytest.mark.parametrize("N", [1, 37, 4097, 70001])
def test_pose_frame_coherent_order_changes_nothing(ctx, player, K, N):
    """pose_sort: the step-ordered kernel visits the instances grouped by (clip, frame) bucket (counting sort on the device).
    Palettes, derived transforms and advanced phases must be bit-identical to the unsorted launch, for pose and for tick
    (where the order is computed from the phase the kernel is about to advance to)."""
    sk, cl, am, mk = player
    c2 = capi_mod().Context(0)  # own context: own clip table, so bucket numbering starts at 0 and stays small enough for pair keys
    try:
        skel = c2.upload_skeleton(sk)
        ids = np.asarray([c2.upload_clip(skel, c, am) for c in (cl, A.make_variant_clip(cl, 17, seed=3), A.make_variant_clip(cl, 29, seed=4))], np.int32)
        rng = np.random.default_rng(N + K)
        pick = ids[rng.integers(0, 3, (N, K))]
        ph = rng.uniform(0, 1, (N, K)).astype(np.float32)
        ph[0, 0] = 0.0
        ph[-1, 0] = np.float32(1.0) - np.float32(2.0 ** -24)
        w = rng.uniform(0.1, 1, (N, K)).astype(np.float32)
        ln = rng.uniform(0.5, 2.0, (N, K)).astype(np.float32)
        out = {}
        c2.set_option("pose_level", 0)  # the property is about the lane-per-instance kernel: same kernel, another visiting order
        for mode in (0, 1):
            c2.set_option("pose_sort", mode)
            b = c2.batch(skel, N, K)
            b.set(clip_ids=pick, phase=ph, weight=w, clip_len=ln)
            b.keep_derived(True)
            b.pose()
            r = [b.download_palette(), b.download_derived()]
            b.tick(1.0 / 30.0, capi_mod().PCL_ADVANCE_ALL)
            b.tick(0.0421, capi_mod().PCL_ADVANCE_FIRST_ONLY)
            r += [b.download_palette(), b.phases()]
            out[mode] = r
            b.close()
        for x, y in zip(out[0], out[1]):
            assert np.array_equal(x, y)
    finally:
        c2.close()


def test_crowd_too_big_for_the_gpu_is_refused_cleanly(ctx, player_dev):
    """A batch whose palettes exceed HBM: PCL_ERR_OOM, nothing leaked, no stale CUDA error at the next launch."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    with pytest.raises(capi.PclError) as e:
        ctx.batch(skel, 70_000_000, 1)  # 70 M x 2880 B = 200 GB of palettes
    assert e.value.code == -4
    mats = gpu_pose(ctx, skel, np.full((64, 1), clip), np.linspace(0, 0.9, 64, dtype=np.float32))
    assert np.isfinite(mats).all()


@pytest.mark.parametrize("K", [1, 2])
def test_pose_register_allocation_variants_agree(ctx, player_dev, oracle, player, K):
    """pose_min_blocks: the step-ordered kernel compiled for 2, 3 and 4 resident blocks per SM (255 / 168 / 128 registers) is the
    same arithmetic, so the palettes must be bit-identical - and right (oracle)."""
    skel, clip = player_dev
    sk, cl, am, mk = player
    N = 3001
    rng = np.random.default_rng(K)
    ph = rng.uniform(0, 1, (N, K)).astype(np.float32)
    w = rng.uniform(0.1, 1, (N, K)).astype(np.float32)
    got = {}
    try:
        for mb in (2, 3, 4):
            ctx.set_option("pose_min_blocks", mb)
            got[mb] = gpu_pose(ctx, skel, np.full((N, K), clip), ph, weight=w)
    finally:
        ctx.set_option("pose_min_blocks", 0)
    assert np.array_equal(got[2], got[3]) and np.array_equal(got[2], got[4])
    osk, ocl = oracle.skeleton(sk), oracle.clip(cl, am)
    ref = oracle.pose_batch(osk, [ocl], np.zeros((N, K), np.int32), ph, weight=w) if K > 1 else \
        oracle.pose_batch(osk, [ocl], np.zeros((N, 1), np.int32), ph[:, 0])
    assert_mats_close(got[2], ref, tol=MAT_TOL if K == 1 else 2e-5)


@pytest.mark.parametrize("K", [1, 2])
def test_set_frame_async_equals_set(ctx, player_dev, K):
    """pcl_batch_set_frame_async: the per-frame inputs through the batch's pinned staging buffer, enqueued without a host
    stall; same device state as pcl_batch_set, for whole batches and sub-ranges, frame after frame."""
    skel, clip = player_dev
    N = 1000
    rng = np.random.default_rng(K)
    a, b = ctx.batch(skel, N, K), ctx.batch(skel, N, K)
    ids = np.full((N, K), clip, np.int32)
    a.set(clip_ids=ids), b.set(clip_ids=ids)
    for frame in range(4):
        ph = rng.uniform(0, 1, (N, K)).astype(np.float32)
        w = rng.uniform(0.1, 1, (N, K)).astype(np.float32)
        a.set(phase=ph, weight=w)
        if frame % 2 == 0:
            b.set_frame_async(phase=ph, weight=w)
        else:  # two sub-ranges, the caller's arrays scribbled over right after the call
            p0, w0 = ph[:333].copy(), w[:333].copy()
            b.set_frame_async(phase=p0, weight=w0, first=0)
            p0[:] = -1.0
            w0[:] = -1.0
            b.set_frame_async(phase=ph[333:], weight=None, first=333)
            b.set_frame_async(phase=None, weight=w[333:], first=333)
        a.pose(), b.pose()
        assert np.array_equal(a.phases(), b.phases())
        assert np.array_equal(a.download_palette(), b.download_palette())
    a.close(), b.close()


@pytest.mark.parametrize("shape", ["player", "chain", "star", "bushy200"])
def test_level_parallel_pose_kernel_matches_the_lane_kernels(ctx, player, oracle, shape):
    """pose_level: small one-clip crowds are posed level by level, 8 warps sharing 32 instances.  Same device functions per bone as
    the lane-per-instance kernels, but compiled in another context, so nvcc contracts a few other multiply-adds: palettes and
    derived transforms agree to a few ulp (2e-6 scaled, five times tighter than the gate against the reference), the advanced
    phases to the bit - for pose and tick, whole and ragged tiles, on a chain (one bone per level), a star (one huge level)
    and a bushy 200-bone tree."""
    from piccolo_b200 import capi
    if shape == "player":
        sk, cl, am, _ = player
    else:
        J, br = {"chain": (40, 1), "star": (90, 89), "bushy200": (200, 3)}[shape]
        sk = A.make_synthetic_skeleton(J, branching=br, seed=11)
        cl, am = A.make_synthetic_clip(sk, 13, seed=12)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    for N in (1, 33, 500):
        ph = np.random.default_rng(N).uniform(0, 1, N).astype(np.float32)
        ln = np.random.default_rng(N + 1).uniform(0.5, 2.0, N).astype(np.float32)
        out = {}
        for mode in (0, 1):
            ctx.set_option("pose_level", mode)
            try:
                b = ctx.batch(skel, N, 1)
                b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph, clip_len=ln)
                b.keep_derived(True)
                b.pose()
                r = [b.download_palette(), b.download_derived()]
                b.tick(1.0 / 30.0, capi.PCL_ADVANCE_FIRST_ONLY)
                b.tick(0.0421, capi.PCL_ADVANCE_ALL)
                r += [b.download_palette(), b.phases()]
                out[mode] = r
                b.close()
            finally:
                ctx.set_option("pose_level", -1)
        assert np.array_equal(out[0][3], out[1][3])  # phases
        assert_mats_close(out[1][0], out[0][0], tol=2e-6)
        assert_mats_close(out[1][2], out[0][2], tol=2e-6)
        d0, d1 = out[0][1], out[1][1]
        assert np.abs(d0 - d1).max() <= 2e-6 * max(1.0, float(np.abs(d0).max()))
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), ph)
        assert_mats_close(out[1][0], ref)


@pytest.mark.parametrize("shape", ["player", "bushy200", "star", "chain"])
def test_two_half_pose_form_is_bit_identical_to_the_whole_walk(ctx, player, oracle, shape):
    """pose_split: two warps per 32 instances, one half of the skeleton each (the cut bone's subtree + its ancestors | the rest).
    Same instantiation of the same device functions, same inputs per bone: every palette row - the ancestors' rows, written by
    both halves, included - equals the whole-skeleton walk's to the bit, on whole and ragged tiles; and the oracle's within the
    gate.  A star (all bones children of the root) and a chain have no balanced cut: the option is then ignored."""
    if shape == "player":
        sk, cl, am, _ = player
    else:
        J, br = {"chain": (40, 1), "star": (90, 89), "bushy200": (200, 3)}[shape]
        sk = A.make_synthetic_skeleton(J, branching=br, seed=11)
        cl, am = A.make_synthetic_clip(sk, 13, seed=12)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    from piccolo_b200 import capi
    for N in (1, 33, 500, 10000):
        ph = np.random.default_rng(N).uniform(0, 1, N).astype(np.float32)
        ln = np.random.default_rng(N + 1).uniform(0.5, 2.0, N).astype(np.float32)
        out = {}
        ctx.set_option("pose_level", 0)
        try:
            for mode in (0, 1):
                ctx.set_option("pose_split", mode)
                b = ctx.batch(skel, N, 1)
                b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph, clip_len=ln)
                b.pose()
                r = [b.download_palette()]
                # ticks: both halves read the old phases, the advanced ones land in a second array the host swaps in
                b.tick(1.0 / 30.0, capi.PCL_ADVANCE_FIRST_ONLY)
                b.tick(0.0421, capi.PCL_ADVANCE_ALL)
                r += [b.download_palette(), b.phases()]
                b.tick(0.0, capi.PCL_ADVANCE_ALL)
                r += [b.download_palette(), b.phases()]
                out[mode] = r
                b.close()
        finally:
            ctx.set_option("pose_level", -1), ctx.set_option("pose_split", -1)
        for x, y in zip(out[0], out[1]):
            assert np.array_equal(x, y)
        assert np.array_equal(out[1][2], out[1][4]) and np.array_equal(out[1][1], out[1][3])  # dt = 0 keeps phases and pose
        ref = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), ph)
        assert_mats_close(out[1][0], ref)
        ref2 = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((N, 1), np.int32), out[1][2])
        assert_mats_close(out[1][1], ref2)


def test_two_contexts_on_one_device_share_kernel_attributes(ctx, player_dev, player):
    """The opt-in shared-memory size is a property of (device, kernel), not of a context: a second context that needs less
    must not shrink what the first one launches with (regression: 'invalid argument' on the first context's next pose)."""
    skel, clip = player_dev
    sk, cl, am, mk = player
    N = 50000
    ph = np.linspace(0, 0.999, N, dtype=np.float32)
    a = gpu_pose(ctx, skel, np.full((N, 1), clip), ph)          # 4 warps per block
    c2 = capi_mod().Context(0)
    try:
        s2 = c2.upload_skeleton(sk)
        k2 = c2.upload_clip(s2, cl, am)
        small = gpu_pose(c2, s2, np.full((40, 1), k2), ph[:40])  # 1 warp per block: a quarter of the shared memory
    finally:
        c2.close()
    b = gpu_pose(ctx, skel, np.full((N, 1), clip), ph)
    assert np.array_equal(a, b)
    assert_mats_close(small, a[:40], tol=2e-6)  # 40 instances take the level-parallel kernel: same values to a few ulp


def test_pose_is_independent_of_batch_composition(ctx, player_dev):
    """Full-size property (config 3 scale): instance i's palette does not depend on N or its position in the batch."""
    skel, clip = player_dev
    N = 100000
    ph = np.modf(np.arange(N, dtype=np.float64) * 0.7548776662)[0].astype(np.float32)
    big = ctx.batch(skel, N, 1)
    big.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
    big.pose()
    sel = np.random.default_rng(0).choice(N, 2000, replace=False)
    ctx.set_option("pose_level", 0)  # same kernel for both crowd sizes: bit-identical; the small-crowd kernel: a few ulp (below)
    try:
        small = gpu_pose(ctx, skel, np.full((2000, 1), clip), ph[sel])
    finally:
        ctx.set_option("pose_level", -1)
    small_level = gpu_pose(ctx, skel, np.full((2000, 1), clip), ph[sel])
    assert_mats_close(small_level, small, tol=2e-6)
    for chunk in range(0, 2000, 500):
        idx = sel[chunk:chunk + 500]
        lo, hi = idx.min(), idx.max() + 1
        blk = big.download_palette(int(lo), int(hi - lo))
        assert np.array_equal(blk[idx - lo], small[chunk:chunk + 500])  # bit-identical
    # idempotence: posing twice changes nothing
    a = big.download_palette(0, 1000)
    big.pose()
    assert np.array_equal(a, big.download_palette(0, 1000))
    big.close()


# ------------------------------------------------------------------------------------------------
# skinning
# ------------------------------------------------------------------------------------------------
def _skin(ctx, skel, clip, mesh_h, V, N, ph, flags, variant, vsplit=0, inst_tile=0, K=1, cta_tile=0):
    ctx.set_option("lbs_variant", variant)
    ctx.set_option("lbs_vsplit", vsplit)
    ctx.set_option("lbs_inst_tile", inst_tile)
    ctx.set_option("lbs_cta_tile", cta_tile)
    try:
        return _skin_body(ctx, skel, clip, mesh_h, V, N, ph, flags)
    finally:
        ctx.set_option("lbs_variant", -1), ctx.set_option("lbs_vsplit", 0), ctx.set_option("lbs_inst_tile", 0)
        ctx.set_option("lbs_cta_tile", 0)


def _skin_body(ctx, skel, clip, mesh_h, V, N, ph, flags):
    b = ctx.batch(skel, N, 1)
    b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
    b.pose()
    sb = ctx.skinbuf(mesh_h, V, N, flags)
    b.skin(mesh_h, sb)
    pos, nt = sb.download()
    pal = b.download_palette()
    sb.close(), b.close()
    return pos, nt, pal


def _check_skin(oracle, mesh, pal, pos, nt):
    for i in range(pal.shape[0]):
        p16 = oracle.palette_slots(pal[i])
        rp, rn, rt = oracle.lbs(mesh, p16)
        radius = max(1.0, float(np.abs(rp).max()))
        assert np.abs(pos[i] - rp).max() <= POS_TOL * radius
        if nt is not None:
            ok = np.isfinite(rn).all(axis=1) & np.isfinite(rt).all(axis=1)
            assert np.abs(nt[i][ok, :3] - rn[ok]).max() <= 1e-4 and np.abs(nt[i][ok, 3:] - rt[ok]).max() <= 1e-4
            assert np.all(np.isnan(nt[i][~ok]).any(axis=1))  # normalize(0) stays NaN like the shader


@pytest.mark.parametrize("variant", [0, 1, 2, 3])
def test_lbs_reference_golden(ctx, player_dev, variant):
    """mesh.vert restated with the reference's Matrix4x4/Vector4 operators (oracle/_ref) on 2048 vertices."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    g = np.load(os.path.join(GOLDEN, "lbs_golden.npz"))
    V = len(g["pos"])
    mesh = A.MeshData(g["pos"], g["nrm"], g["tan"], np.zeros((V, 2), np.float32), np.arange(V, dtype=np.int32), g["idx"], g["w"])
    mh = ctx.upload_mesh(mesh, 58)
    pos, nt, pal = _skin(ctx, skel, clip, mh, V, 1, np.float32([0.37]), capi.PCL_SKIN_POS_NRM_TAN, variant)
    assert_mats_close(pal[0], g["palette"][1:])
    assert np.abs(pos[0] - g["out_pos"]).max() <= POS_TOL * max(1.0, np.abs(g["out_pos"]).max())
    ok = np.isfinite(g["out_nrm"]).all(axis=1)
    assert np.abs(nt[0][ok, :3] - g["out_nrm"][ok]).max() <= 1e-4 and np.abs(nt[0][ok, 3:] - g["out_tan"][ok]).max() <= 1e-4
    assert np.all(pos[0][5] == 0) and np.all(np.isnan(nt[0][5]))  # vertex with zero total weight


@pytest.mark.parametrize("variant", [0, 1, 2, 3])
@pytest.mark.parametrize("V,N", [(16380, 37), (1001, 5), (255, 3), (4, 1), (513, 70)])
def test_lbs_vs_oracle_player(ctx, player_dev, oracle, player, variant, V, N):
    from piccolo_b200 import capi
    skel, clip = player_dev
    mesh = A.make_player_mesh(player[0], V=max(V, 58), seed=42)
    if V < 58:
        mesh = A.MeshData(*(a[:V].copy() for a in (mesh.pos, mesh.nrm, mesh.tan, mesh.uv, mesh.index_buffer, mesh.bind_idx, mesh.bind_w)))
    mh = ctx.upload_mesh(mesh, 58)
    ph = np.random.default_rng(V).uniform(0, 1, N).astype(np.float32)
    for flags in (capi.PCL_SKIN_POS_NRM_TAN, capi.PCL_SKIN_POS):
        for vsplit, it, warps in ((0, 0, 0), (1, 16, 8), (3, 5, 12)):
            ctx.set_option("lbs_warps", warps)
            pos, nt, pal = _skin(ctx, skel, clip, mh, V, N, ph, flags, variant, vsplit, it)
            _check_skin(oracle, mesh, pal, pos, nt)
    ctx.set_option("lbs_warps", 0)


@pytest.mark.parametrize("variant", [0, 1, -1, 2, 3])
def test_lbs_config5_synthetic(ctx, oracle, variant):
    """257 palette slots: 32 palettes do not fit in shared memory.  The TMA-store crowd kernel (3) drops to 8 instances per
    CTA (4 vertex streams per warp; the ragged last tile of 3 instances is clipped by the launch's tensor maps); forcing the
    LSU crowd kernel (2), which only exists for 32 palettes, must be refused, not silently mis-skinned."""
    from piccolo_b200 import capi
    sk = A.make_synthetic_skeleton(256)
    cl, am = A.make_synthetic_clip(sk, 32)
    mesh = A.make_synthetic_mesh(sk, V=49998)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    mh = ctx.upload_mesh(mesh, 256)
    N = 43
    ph = np.linspace(0.05, 0.95, N, dtype=np.float32)
    if variant == 2:
        with pytest.raises(capi.PclError) as e:
            _skin(ctx, skel, clip, mh, mesh.V, N, ph, capi.PCL_SKIN_POS_NRM_TAN, variant)
        assert e.value.code == -3
        ctx.set_option("lbs_variant", -1)
        return
    pos, nt, pal = _skin(ctx, skel, clip, mh, mesh.V, N, ph, capi.PCL_SKIN_POS_NRM_TAN, variant)
    _check_skin(oracle, mesh, pal[::5], pos[::5], nt[::5])


@pytest.mark.parametrize("J,variant", [(75, 3), (75, 2), (82, 3), (83, 2), (83, -1), (60, 3), (83, 3), (120, 3), (165, 3),
                                       (166, 3), (200, -1), (340, 3), (400, 3)])
def test_lbs_crowd_kernels_on_mid_sized_skeletons(ctx, oracle, J, variant):
    """The TMA-store crowd kernel keeps 32 palettes in shared memory up to 82 bones (3 staging tiles up to 63), 16 palettes
    up to ~165 bones and 8 up to ~340; the LSU-flush kernel exists for 32 palettes only.  Beyond their limits they must
    refuse (explicit variant) or fall back (auto)."""
    from piccolo_b200 import capi
    sk = A.make_synthetic_skeleton(J, branching=2, seed=J)
    cl, am = A.make_synthetic_clip(sk, 12, seed=J + 1)
    mesh = A.make_player_mesh(sk, V=3000, seed=J)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    mh = ctx.upload_mesh(mesh, J)
    N = 192 + (J % 7)  # ragged against 32 / 16 / 8 for some J
    ph = np.random.default_rng(J).uniform(0, 1, N).astype(np.float32)
    if (J > 82 and variant == 2) or J >= 400:
        ctx.set_option("lbs_compact", 0)  # whole palettes only: 400 bones x 8 instances do not fit
        try:
            with pytest.raises(capi.PclError) as e:
                _skin(ctx, skel, clip, mh, mesh.V, N, ph, capi.PCL_SKIN_POS_NRM_TAN, variant)
            assert e.value.code == -3
        finally:
            ctx.set_option("lbs_variant", -1), ctx.set_option("lbs_compact", -1)
        if variant == 2:
            return  # (variant 3 goes on: the compact form below takes a coherent 400-bone mesh)
    pos, nt, pal = _skin(ctx, skel, clip, mh, mesh.V, N, ph, capi.PCL_SKIN_POS_NRM_TAN, variant)
    _check_skin(oracle, mesh, pal[::37], pos[::37], nt[::37])


@pytest.mark.parametrize("J,V,N,compact", [(58, 4100, 70, 1), (120, 9000, 45, 1), (256, 20000, 37, -1), (256, 20000, 64, 1), (400, 6000, 33, -1)])
@pytest.mark.parametrize("flags_nt", [True, False])
def test_lbs_compact_form_block_local_palette_rows(ctx, oracle, J, V, N, compact, flags_nt):
    """lbs_cta_kernel's compact form: per block of 1024 vertices the CTA keeps only the palette rows that block references
    and the records carry block-local row numbers.  Forced on skeletons whose whole palettes would fit too (player-sized, 120
    bones), chosen automatically for coherent 256- and 400-bone characters; ragged instance counts, vertex counts that are
    not a multiple of a block, both output sets; every result against the oracle and against the whole-palette kernels."""
    from piccolo_b200 import capi
    sk = A.make_synthetic_skeleton(J, branching=2, seed=J) if J != 58 else A.load_player()[0]
    cl, am = (A.make_synthetic_clip(sk, 12, seed=J + 1) if J != 58 else A.load_player()[1:3])
    mesh = A.make_coherent_mesh(sk, V=V, seed=J)
    skel = ctx.upload_skeleton(sk)
    clip = ctx.upload_clip(skel, cl, am)
    mh = ctx.upload_mesh(mesh, J)
    ph = np.random.default_rng(J + N).uniform(0, 1, N).astype(np.float32)
    flags = capi.PCL_SKIN_POS_NRM_TAN if flags_nt else capi.PCL_SKIN_POS
    ctx.set_option("lbs_compact", compact)
    try:
        pos, nt, pal = _skin(ctx, skel, clip, mh, mesh.V, N, ph, flags, 3)
    finally:
        ctx.set_option("lbs_compact", -1)
    sel = sorted({0, N - 1, N // 2, N // 3})
    _check_skin(oracle, mesh, pal[sel], pos[sel], nt[sel] if nt is not None else None)
    if J <= 340:  # the whole-palette form of the same kernel (another tile size for the big skeletons): same sums, slot order may differ
        ctx.set_option("lbs_compact", 0)
        try:
            pos0, nt0, _ = _skin(ctx, skel, clip, mh, mesh.V, N, ph, flags, 3)
        finally:
            ctx.set_option("lbs_compact", -1)
        assert np.abs(pos0 - pos).max() <= 1e-5 * max(1.0, float(np.abs(pos0).max()))


@pytest.mark.parametrize("cta_tile", [16, 8])
@pytest.mark.parametrize("flags_nt", [True, False])
def test_lbs_tma_store_kernel_smaller_instance_tiles_on_player(ctx, player_dev, oracle, player, cta_tile, flags_nt):
    """16 and 8 instances per CTA (2 and 4 vertex streams per warp) forced on the player, whose 32 palettes would fit:
    same results as the oracle, and the 32-instance form's up to the order of the four influence products (the slot packing
    follows each form's own visiting order)."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    mesh = A.make_player_mesh(player[0], V=5000, seed=7)
    mh = ctx.upload_mesh(mesh, 58)
    N = 208  # a multiple of 16 and 8: no remainder kernel involved
    ph = np.random.default_rng(cta_tile).uniform(0, 1, N).astype(np.float32)
    flags = capi.PCL_SKIN_POS_NRM_TAN if flags_nt else capi.PCL_SKIN_POS
    pos, nt, pal = _skin(ctx, skel, clip, mh, mesh.V, N, ph, flags, 3, cta_tile=cta_tile)
    _check_skin(oracle, mesh, pal[::41], pos[::41], None if nt is None else nt[::41])
    pos32, nt32, _ = _skin(ctx, skel, clip, mh, mesh.V, N - 16, ph[:N - 16], flags, 3, cta_tile=32)
    assert np.abs(pos[:N - 16] - pos32).max() <= 4e-6 * max(1.0, float(np.abs(pos32).max()))
    if flags_nt:
        ok = np.isfinite(nt32)
        assert np.array_equal(ok, np.isfinite(nt[:N - 16])) and np.abs(nt[:N - 16][ok] - nt32[ok]).max() <= 4e-6


@pytest.mark.parametrize("cta_tile", [32, 16, 8])
def test_lbs_ragged_last_tile_is_clipped_by_the_tensor_maps(ctx, player_dev, player, cta_tile):
    """An instance count that is not a multiple of the tile: the last tile's missing instances are dropped by tensor maps encoded
    for this launch (rows = slot_first + count).  The ring rows behind the launch must keep what was there, bit for bit, and the
    rows of the launch must equal the lane-per-vertex kernel's."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    mesh = A.make_player_mesh(player[0], V=3000, seed=11)
    mh = ctx.upload_mesh(mesh, 58)
    cap, N, slot0 = 260, 197, 7  # 197 = 6 x 32 + 5 = 12 x 16 + 5 = 24 x 8 + 5
    rng = np.random.default_rng(cta_tile)
    sb = ctx.skinbuf(mh, mesh.V, cap, capi.PCL_SKIN_POS_NRM_TAN)
    old = ctx.batch(skel, cap, 1)
    old.set(clip_ids=np.full((cap, 1), clip, np.int32), phase=rng.uniform(0, 1, cap).astype(np.float32))
    old.pose()
    old.skin(mh, sb)
    before_pos, before_nt = sb.download()
    new = ctx.batch(skel, N, 1)
    new.set(clip_ids=np.full((N, 1), clip, np.int32), phase=rng.uniform(0, 1, N).astype(np.float32))
    new.pose()
    try:
        ctx.set_option("lbs_variant", 3), ctx.set_option("lbs_cta_tile", cta_tile)
        new.skin(mh, sb, 0, N, slot0)
        pos, nt = sb.download()
        ctx.set_option("lbs_variant", 0)
        new.skin(mh, sb, 0, N, slot0)
        ref_pos, ref_nt = sb.download()
    finally:
        ctx.set_option("lbs_variant", -1), ctx.set_option("lbs_cta_tile", 0)
    keep = np.ones(cap, bool)
    keep[slot0:slot0 + N] = False
    assert np.array_equal(pos[keep], before_pos[keep]) and np.array_equal(nt[keep], before_nt[keep], equal_nan=True)
    assert np.abs(pos[~keep] - ref_pos[~keep]).max() <= 4e-6 * max(1.0, float(np.abs(ref_pos).max()))
    ok = np.isfinite(ref_nt[~keep])
    assert np.abs(nt[~keep][ok] - ref_nt[~keep][ok]).max() <= 4e-6
    for x in (old, new):
        x.close()
    sb.close()


@pytest.mark.parametrize("variant", [0, 2, 3])
def test_lbs_ring_slots_and_linearity(ctx, player_dev, player, variant):
    """Ring addressing + a size-independent property: skinning is linear in the vertex position stream."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    mesh = A.make_player_mesh(player[0], V=16380, seed=1)
    m2 = A.MeshData(mesh.pos * np.float32(2.0), mesh.nrm, mesh.tan, mesh.uv, mesh.index_buffer, mesh.bind_idx, mesh.bind_w)
    h1, h2 = ctx.upload_mesh(mesh, 58), ctx.upload_mesh(m2, 58)
    N, C = 160, 32
    b = ctx.batch(skel, N, 1)
    b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=np.linspace(0, 1, N, dtype=np.float32))
    b.pose()
    ctx.set_option("lbs_variant", variant)
    s1, s2 = ctx.skinbuf(h1, 16380, 3 * C, capi.PCL_SKIN_POS), ctx.skinbuf(h2, 16380, 3 * C, capi.PCL_SKIN_POS)
    b.skin(h1, s1, first=96, count=C, slot_first=C)
    b.skin(h2, s2, first=96, count=C, slot_first=C)
    b.skin(h1, s1, first=0, count=4, slot_first=0)
    p1, _ = s1.download(C, C)
    p2, _ = s2.download(C, C)
    pal = b.download_palette(96, C)
    t = pal[:, :, [3, 7, 11]]  # translations; p(2x) - 2 p(x) = -sum_k w_k t_k : bounded by max |t|
    assert np.abs(p2 - 2 * p1).max() <= np.abs(t).max() * 1.0001 + 1e-4
    full = ctx.skinbuf(h1, 16380, N, capi.PCL_SKIN_POS)
    b.skin(h1, full)
    pf, _ = full.download()
    ctx.set_option("lbs_vsplit", 5)
    b.skin(h1, s1, first=96, count=C, slot_first=2 * C)
    ctx.set_option("lbs_vsplit", 0), ctx.set_option("lbs_variant", -1)
    assert np.array_equal(pf[96:96 + C], p1)  # same instances through a different slot / tile position: bit-identical
    assert np.array_equal(s1.download(2 * C, C)[0], p1)  # and through a different vertex split
    small = s1.download(0, 4)[0]  # 4 instances: variant 3 hands them to the variant-2 kernel (another summation order)
    assert np.array_equal(pf[0:4], small) if variant != 3 else np.abs(pf[0:4] - small).max() <= 1e-5
    for x_ in (s1, s2, full, b):
        x_.close()


@pytest.mark.parametrize("variant", [2, 3])
def test_lbs_full_size_config3_properties(ctx, oracle, player, variant):
    """BASELINE.json config 3 at full size (100 000 instances, 2-clip blend, 16 380 vertices, 25 000-slot ring): sampled
    instances match the oracle, and an instance's skinned streams do not depend on the launch it was part of."""
    from piccolo_b200 import capi
    sk, cl, am, mk = player
    run = A.make_variant_clip(cl, 17, seed=1234)
    mesh = A.make_player_mesh(sk, V=16380, seed=42)
    skel = ctx.upload_skeleton(sk)
    c0, c1 = ctx.upload_clip(skel, cl, am), ctx.upload_clip(skel, run, am)
    mh = ctx.upload_mesh(mesh, 58)
    N, ring = 100000, 25000
    i = np.arange(N, dtype=np.float64)
    w = np.modf(i * 0.7548776662)[0].astype(np.float32)
    weight = np.stack([w, np.float32(1) - w], 1)
    phase = np.stack([np.modf(i * 0.6180339887)[0], np.modf(i * 0.6180339887 + 0.5)[0]], 1).astype(np.float32)
    b = ctx.batch(skel, N, 2)
    b.set(clip_ids=np.tile(np.int32([c0, c1]), (N, 1)), phase=phase, weight=weight)
    b.pose()
    sb = ctx.skinbuf(mh, mesh.V, ring, capi.PCL_SKIN_POS_NRM_TAN)
    small = ctx.skinbuf(mh, mesh.V, 64, capi.PCL_SKIN_POS_NRM_TAN)
    ctx.set_option("lbs_variant", variant)
    orc_clips = [oracle.clip(cl, am), oracle.clip(run, am)]
    sv = oracle.skeleton(sk)
    for chunk in (0, 3):
        b.skin(mh, sb, first=chunk * ring, count=ring, slot_first=0)
        for slot in (0, 12345, ring - 1):
            inst = chunk * ring + slot
            pos, nt = sb.download(slot, 1)
            ref = oracle.pose_batch(sv, orc_clips, np.int32([[0, 1]]), phase[inst:inst + 1], weight=weight[inst:inst + 1])
            rp, rn, rt = oracle.lbs(mesh, oracle.palette_slots(ref[0]))
            assert np.abs(pos[0] - rp).max() <= POS_TOL * max(1.0, np.abs(rp).max())
            assert np.abs(nt[0][:, :3] - rn).max() <= 1e-4 and np.abs(nt[0][:, 3:] - rt).max() <= 1e-4
            lo = min(max(inst // 32 * 32 - 32, 0), N - 64)  # the same instance inside another 64-instance launch ...
            if variant != 3 and lo + 67 <= N and inst >= lo + 3:
                lo += 3  # ... and, where the kernel allows ragged tiles, in another lane of its tile
            b.skin(mh, small, first=lo, count=64, slot_first=0)
            p2, n2 = small.download(inst - lo, 1)
            if variant == 3 and slot >= ring // 32 * 32:
                # the ragged tail of the 25 000-instance launch was skinned by the variant-2 kernel (other summation order)
                assert np.abs(p2 - pos).max() <= 1e-5 and np.abs(n2 - nt).max() <= 1e-5
            else:
                assert np.array_equal(p2, pos) and np.array_equal(n2, nt)
    ctx.set_option("lbs_variant", -1)
    for x in (sb, small, b):
        x.close()


def test_async_palette_download_overlapping_skin_and_next_pose(ctx, player_dev, player):
    """The end-to-end path of bench.py: palette download on the copy stream into pinned memory while the compute stream
    skins, then the next pose must wait for the export to have read the palette."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    mesh = A.make_player_mesh(player[0], V=4096, seed=2)
    mh = ctx.upload_mesh(mesh, 58)
    N = 5000
    b = ctx.batch(skel, N, 1)
    sb = ctx.skinbuf(mh, mesh.V, N, capi.PCL_SKIN_POS)
    pinned = ctx.pinned((N, 58, 16), np.float32)
    ph = np.random.default_rng(9).uniform(0, 1, (3, N)).astype(np.float32)
    want = []
    for r in range(3):  # reference results, fully synchronous
        b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph[r])
        b.pose()
        want.append(b.download_palette())
    for r in range(3):
        b.set(phase=ph[r])
        b.pose()
        b.download_palette(out=pinned, asynchronous=True)
        b.skin(mh, sb)
        b.set(phase=ph[(r + 1) % 3])  # next frame's inputs and pose are enqueued before the download has been waited for
        b.pose()
        ctx.sync()
        assert np.array_equal(pinned, want[r]), "async download saw the next frame's palette (round %d)" % r
    sb.close(), b.close()


@pytest.mark.parametrize("pinned_dst", [True, False])
@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_compact_3x4_download_all_modes(ctx, player_dev, player, mode, pinned_dst):
    """pcl_batch_download_palette3x4(_async): rows 0..2 of every AnimationResult matrix, bit-identical to the 4x4 download
    (whose row 3 is the constant 0 0 0 1), for every way the bytes can travel ("export_mode"), into pinned and pageable
    memory, for sub-ranges, and overlapped with the skinning and the NEXT frame's pose like bench.py's end-to-end loop."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    mesh = A.make_player_mesh(player[0], V=4096, seed=2)
    mh = ctx.upload_mesh(mesh, 58)
    N = 5000
    b = ctx.batch(skel, N, 1)
    sb = ctx.skinbuf(mh, mesh.V, N, capi.PCL_SKIN_POS)
    dst = ctx.pinned((N, 58, 12), np.float32) if pinned_dst else np.zeros((N, 58, 12), np.float32)
    ph = np.random.default_rng(19).uniform(0, 1, (3, N)).astype(np.float32)
    ctx.set_option("export_mode", mode)
    try:
        want = []
        for r in range(3):
            b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph[r])
            b.pose()
            want.append(b.download_palette())
            assert np.array_equal(want[r][..., 12:], np.broadcast_to(np.float32([0, 0, 0, 1]), want[r][..., 12:].shape))
            assert np.array_equal(capi.expand_3x4(b.download_palette3x4()), want[r])  # synchronous form
            sub = b.download_palette3x4(first=1234, count=77)
            assert np.array_equal(sub, want[r][1234:1311, :, :12])
        for r in range(3):
            b.set(phase=ph[r])
            b.pose()
            dst[...] = -1.0
            b.download_palette3x4(out=dst, asynchronous=True)
            b.skin(mh, sb)
            b.set(phase=ph[(r + 1) % 3])  # the next frame is enqueued before the download has been waited for
            b.pose()
            ctx.sync()
            assert np.array_equal(dst, want[r][..., :12]), "3x4 download saw the next frame's palette (mode %d, round %d)" % (mode, r)
    finally:
        ctx.set_option("export_mode", 0)
        sb.close(), b.close()


@pytest.mark.parametrize("N,K", [(1, 1), (8, 1), (100, 1), (1000, 1), (5, 2)])
def test_one_call_frame_for_small_batches(ctx, player, oracle, N, K):
    """pcl_batch_frame: inputs + pose + AnimationResult rows in one call (one launch, one sync) for a handful of characters.
    Rows match the oracle and the regular set / pose / download path; NULL inputs keep the previous frame's; the device
    palette is the one a following pcl_batch_skin reads; the asynchronous form returns the same rows later."""
    from piccolo_b200 import capi
    sk, cl, am, mk = player
    clips = [cl, A.make_variant_clip(cl, 17, seed=1234)]
    skel = ctx.upload_skeleton(sk)
    ids = np.int32([ctx.upload_clip(skel, c, am) for c in clips])
    rng = np.random.default_rng(100 + N)
    pick = rng.integers(0, 2, (N, K)).astype(np.int32)
    if K == 2:
        pick[:, 1] = 1 - pick[:, 0]
    ph = rng.uniform(0, 1, (2, N, K)).astype(np.float32)
    w = rng.uniform(0.2, 1, (N, K)).astype(np.float32)
    b, b2 = ctx.batch(skel, N, K), ctx.batch(skel, N, K)
    try:
        sv, ocl = oracle.skeleton(sk), [oracle.clip(c, am) for c in clips]
        l0 = ctx.launch_count()
        rows = b.frame(clip_ids=ids[pick], phase=ph[0], weight=w if K > 1 else None)
        assert ctx.launch_count() - l0 == (2 if (K > 1 and N <= 8) else 1)  # K = 1: the pose kernel writes the rows itself or a DMA fetches them
        ref = oracle.pose_batch(sv, ocl, pick, ph[0], weight=w if K > 1 else None)
        assert_mats_close(capi.expand_3x4(rows), ref, tol=MAT_TOL if K == 1 else 2e-5)
        b2.set(clip_ids=ids[pick], phase=ph[0], weight=w if K > 1 else None)
        b2.pose()
        assert np.abs(b2.download_palette3x4() - rows).max() <= 2e-6
        assert np.array_equal(b.download_palette3x4(), rows)  # the device palette is this frame's
        # second frame: only the phases change; asynchronous form
        assert b.frame(phase=ph[1], wait=False) is None
        rows2 = b.frame_result()
        ref2 = oracle.pose_batch(sv, ocl, pick, ph[1], weight=w if K > 1 else None)
        assert_mats_close(capi.expand_3x4(rows2), ref2, tol=MAT_TOL if K == 1 else 2e-5)
        with pytest.raises(capi.PclError):
            b.frame(clip_ids=np.full((N, K), 10 ** 6, np.int32), phase=ph[0])
    finally:
        b.close(), b2.close()


def test_one_call_frame_refuses_large_batches(ctx, player_dev):
    from piccolo_b200 import capi
    skel, clip = player_dev
    N = 30000  # 30 000 x 58 x 48 B > PCL_FRAME_MAX_BYTES
    b = ctx.batch(skel, N, 1)
    try:
        with pytest.raises(capi.PclError) as e:
            b.frame(clip_ids=np.full((N, 1), clip, np.int32), phase=np.zeros((N, 1), np.float32))
        assert e.value.code == -3
    finally:
        b.close()


def test_released_assets_give_memory_and_buckets_back(player):
    """Level reloads: clips, meshes and skeletons can be released (the reference's AnimationManager caches only ever grow,
    animation_system.cpp:10-13).  400 upload / release cycles leave the device memory where it was and the (clip, frame)
    bucket count low enough for the frame-coherent order (each bundled clip occupies 23 buckets, the order supports 8192);
    a released handle is refused; a skeleton in use cannot be released."""
    import torch
    from piccolo_b200 import capi
    sk, cl, am, mk = player
    mesh = A.make_player_mesh(sk, V=3000, seed=5)
    c = capi.Context(0)
    try:
        skel = c.upload_skeleton(sk)
        keep = c.upload_clip(skel, cl, am)
        b = c.batch(skel, 5000, 1)
        ph = np.linspace(0, 1, 5000, endpoint=False, dtype=np.float32)
        b.set(clip_ids=np.full((5000, 1), keep, np.int32), phase=ph)
        b.pose()
        want = b.download_palette3x4()
        torch.cuda.init()
        c.sync()
        for it in range(400):
            if it == 20:
                free0 = torch.cuda.mem_get_info(0)[0]  # (the allocator's pools have settled)
            tmp = c.upload_clip(skel, cl, am)
            m = c.upload_mesh(mesh, sk.J)
            s2 = c.upload_skeleton(sk)
            c.release_clip(tmp), c.release_mesh(m), c.release_skeleton(s2)
        assert torch.cuda.mem_get_info(0)[0] >= free0 - (4 << 20)
        with pytest.raises(capi.PclError):
            c.release_clip(tmp)  # twice
        with pytest.raises(capi.PclError):
            b.set(clip_ids=np.full((5000, 1), tmp, np.int32), phase=ph)  # a released clip
        with pytest.raises(capi.PclError):
            c.release_skeleton(skel)  # a batch and a clip still use it
        c.set_option("pose_sort", 1)
        l0 = c.launch_count()
        b.set(clip_ids=np.full((5000, 1), keep, np.int32), phase=ph)
        b.pose()
        assert c.launch_count() - l0 == 4, "the frame-coherent order was lost to leaked buckets"
        assert np.array_equal(b.download_palette3x4(), want)
        b.close()
        c.release_clip(keep)
        c.release_skeleton(skel)
    finally:
        c.close()


def test_render_handoff_layout(ctx, player_dev):
    """joint_matrices[1024 * instance_in_draw + slot], 64 instances per draw, slot 0 = identity, only J+1 slots touched
    (main_camera_pass.cpp:2284-2298; mesh_component.cpp:72-79)."""
    from piccolo_b200 import capi
    skel, clip = player_dev
    N = 150
    b = ctx.batch(skel, N, 1)
    b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=np.linspace(0, 1, N, endpoint=False, dtype=np.float32))
    b.pose()
    pal = b.download_palette()
    rb = capi.RenderBuf(ctx, 4)
    b.export_render(rb, first=0, count=N, draw_first=1)  # draws 1..3
    eye = np.eye(4, dtype=np.float32).reshape(16)
    assert not rb.download(0).any()
    for d in range(3):
        m = rb.download(1 + d)
        n_in = min(64, N - 64 * d)
        assert np.array_equal(m[:n_in, 1:59], pal[64 * d:64 * d + n_in])  # bit-identical to AnimationResult
        assert np.array_equal(m[:n_in, 0], np.broadcast_to(eye, (n_in, 16)))
        assert not m[:n_in, 59:].any() and not m[n_in:].any()
    with pytest.raises(capi.PclError):
        b.export_render(rb, first=0, count=N, draw_first=2)  # would need draws 2..4
    rb.close(), b.close()


# ------------------------------------------------------------------------------------------------
# boundary behaviour
# ------------------------------------------------------------------------------------------------
def test_upload_errors(ctx, player):
    from piccolo_b200 import capi
    sk, cl, am, mk = A.load_player()
    sk.is_flat = False
    with pytest.raises(capi.PclError) as e:
        ctx.upload_skeleton(sk)
    assert e.value.code == -3  # the reference silently builds nothing; the C ABI says so
    sk.is_flat = True
    sk.tpose[3, 12] = 0.5
    with pytest.raises(capi.PclError):
        ctx.upload_skeleton(sk)
    sk, cl, am, mk = A.load_player()
    skel = ctx.upload_skeleton(sk)
    bad = A.AnimSkelMap(am.convert.copy())
    bad.convert[3] = 58
    with pytest.raises(capi.PclError):
        ctx.upload_clip(skel, cl, bad)
    cl.rot_keys[7] = cl.rot_keys[7][:0]
    with pytest.raises(capi.PclError):
        ctx.upload_clip(skel, cl, am)
    sk, cl, am, mk = A.load_player()
    cl.total_frame = 0  # would shrink the frame buckets of the coherent order and index out of them
    with pytest.raises(capi.PclError) as e:
        ctx.upload_clip(skel, cl, am)
    assert e.value.code == -1 and "total_frame" in str(e.value)
    import ctypes as C
    counts = np.ones((1, 3), np.int32)  # one channel with one key per track, but no key arrays
    conv = np.zeros(1, np.int32)
    out = C.c_int32()
    rc = ctx.L.pcl_upload_clip(ctx.h, skel, 2, 1, 1, counts.ctypes.data_as(C.POINTER(C.c_int32)), None, None, None, 1,
                               conv.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(out))
    assert rc == -1
    mesh = A.make_player_mesh(sk, V=100)
    mesh.bind_idx[0, 0] = 59
    with pytest.raises(capi.PclError):
        ctx.upload_mesh(mesh, 58)
    with pytest.raises(capi.PclError):
        ctx.batch(skel, 10, 5)
    b = ctx.batch(skel, 10, 1)
    with pytest.raises(capi.PclError):
        b.download_palette()  # nothing posed yet
    b.close()


# ------------------------------------------------------------------------------------------------
# palette pool: palettes of several batches gathered in one buffer
# ------------------------------------------------------------------------------------------------
def test_palette_pool_gathers_shards_for_one_consumer(ctx, player_dev, player, oracle):
    """Two producer batches (the shards of a crowd) pose straight into one pool; a consumer batch attached to all rows adopts
    them and downloads / skins exactly what one big batch produces.  On one GPU this exercises everything but the IPC mapping
    (tools/pool_gather_check.py does that under torchrun on two GPUs)."""
    from piccolo_b200 import capi
    from piccolo_b200.sharding import shard_range
    skel, clip = player_dev
    N = 777
    ph = np.random.default_rng(5).uniform(0, 1, N).astype(np.float32)
    whole = ctx.batch(skel, N, 1)
    whole.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
    whole.pose()
    ref = whole.download_palette()
    mesh = A.make_player_mesh(player[0], V=2000, seed=3)
    mh = ctx.upload_mesh(mesh, 58)
    sb = ctx.skinbuf(mh, mesh.V, N, capi.PCL_SKIN_POS_NRM_TAN)
    whole.skin(mh, sb)
    ref_pos, ref_nt = sb.download()

    pool = ctx.pool(skel, N)
    shards = []
    for r in range(2):
        first, count = shard_range(N, 2, r)
        b = ctx.batch(skel, count, 1)
        b.attach_pool(pool, first)
        b.set(clip_ids=np.full((count, 1), clip, np.int32), phase=ph[first:first + count])
        b.pose()
        assert np.array_equal(b.download_palette(), ref[first:first + count])  # a producer can still read its own rows
        shards.append(b)
    ctx.sync()
    consumer = ctx.batch(skel, N, 1)
    consumer.attach_pool(pool, 0)
    with pytest.raises(capi.PclError):
        consumer.download_palette()  # nothing adopted yet
    consumer.adopt_palettes()
    assert np.array_equal(consumer.download_palette(), ref)
    consumer.skin(mh, sb)
    pos, nt = sb.download()
    assert np.array_equal(pos, ref_pos) and np.array_equal(nt, ref_nt, equal_nan=True)
    # the copy-engine way of filling the same rows: an unattached batch pushes its palettes
    pool2 = ctx.pool(skel, N + 3)
    whole.push_to_pool(pool2, 3)
    ctx.sync()
    view = ctx.batch(skel, N, 1)
    view.attach_pool(pool2, 3)
    view.adopt_palettes()
    assert np.array_equal(view.download_palette(), ref)
    view.close()
    pool2.close()
    with pytest.raises(capi.PclError):
        shards[0].push_to_pool(pool, 0)         # attached batches have nothing to push
    # misuse is refused, not undefined
    with pytest.raises(capi.PclError):
        consumer.attach_pool(pool, 0)           # already attached
    with pytest.raises(capi.PclError):
        pool.close()                            # batches still attached
    late = ctx.batch(skel, 10, 1)
    with pytest.raises(capi.PclError):
        late.attach_pool(pool, N - 5)           # rows outside the pool
    other = ctx.upload_skeleton(A.make_synthetic_skeleton(20))
    wrong = ctx.batch(other, 4, 1)
    with pytest.raises(capi.PclError):
        wrong.attach_pool(pool, 0)              # another skeleton's row layout
    for b in shards + [consumer, late, wrong, whole]:
        b.close()
    sb.close()
    pool.close()


def test_palette_pool_across_two_gpus_under_torchrun(tmp_path):
    """The real thing: rank 1's pose kernel stores its shard's palettes into rank 0's memory over NVLink (CUDA IPC mapping);
    needs two GPUs, so the single-GPU round-end run skips it (run by hand with `gpurun --gpus 2`)."""
    import socket, subprocess, sys, json
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "pool_gather_check.py"), "--instances", "20000"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["identical"] is True and line["world"] == 2


def test_two_contexts_driven_from_two_threads(player):
    """Two contexts on one GPU, each driven by its own host thread at the same time (ctypes releases the GIL in every call):
    different skeletons, so the shared constant bank of the step-ordered pose kernel changes hands on every launch; the results
    must be those of the sequential runs."""
    import threading
    from piccolo_b200 import capi
    sk, cl, am, mk = player
    sk2 = A.make_synthetic_skeleton(40, branching=3, seed=4)
    cl2, am2 = A.make_synthetic_clip(sk2, 11, seed=5)
    jobs = [(sk, cl, am, 3001), (sk2, cl2, am2, 2500)]

    def run(job, rounds, out):
        s, c, m, n = job
        ctx = capi.Context(0)
        try:
            skel = ctx.upload_skeleton(s)
            clip = ctx.upload_clip(skel, c, m)
            mesh = A.make_player_mesh(s, V=1500, seed=n)
            mh = ctx.upload_mesh(mesh, s.J)
            b = ctx.batch(skel, n, 1)
            sb = ctx.skinbuf(mh, mesh.V, n, capi.PCL_SKIN_POS_NRM_TAN)
            b.set(clip_ids=np.full((n, 1), clip, np.int32), phase=np.linspace(0, 0.95, n, dtype=np.float32))
            for _ in range(rounds):
                b.tick(1.0 / 60.0, capi.PCL_ADVANCE_ALL)
                b.skin(mh, sb)
            out.append((b.download_palette(), sb.download()[0]))
            sb.close(), b.close()
        except Exception as e:  # surfaced by the assertion below
            out.append(e)
        finally:
            ctx.close()

    want = []
    for j in jobs:
        o = []
        run(j, 25, o)
        want.append(o[0])
    got = [[], []]
    th = [threading.Thread(target=run, args=(jobs[i], 25, got[i])) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for i in range(2):
        assert not isinstance(got[i][0], Exception), got[i][0]
        assert np.array_equal(got[i][0][0], want[i][0]) and np.array_equal(got[i][0][1], want[i][1])
