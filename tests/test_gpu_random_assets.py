"""Randomised assets against the oracle: tree shapes (forests, stars, chains, bushy), non-unit bind quaternions,
non-uniform scales, arbitrary affine inverse binds, ragged key tracks (clamp carry-over), skipped and duplicated
channel->bone entries, node_count / convert length mismatches, K = 1 and blended K = 2 with masks."""
import numpy as np
import pytest

from piccolo_b200 import assets as A

pytestmark = pytest.mark.gpu


def random_assets(seed):
    rng = np.random.default_rng(seed)
    J = int(rng.integers(2, 90))
    shape = seed % 4
    parent = np.zeros(J, np.int32)
    parent[0] = A.INT_MAX
    for i in range(1, J):
        if shape == 0:
            parent[i] = rng.integers(0, i)                      # bushy
        elif shape == 1:
            parent[i] = i - 1 if rng.uniform() < 0.9 else rng.integers(0, i)  # mostly chains
        elif shape == 2:
            parent[i] = 0 if rng.uniform() < 0.7 else rng.integers(0, i)      # star-like
        else:
            parent[i] = A.INT_MAX if rng.uniform() < 0.1 else (i + 3 if rng.uniform() < 0.05 else rng.integers(0, i))  # forest; parent >= i => root
    q = rng.normal(size=(J, 4)) * rng.uniform(0.2, 3.0, size=(J, 1))   # not normalised: Bone::initialize normalises
    inv = np.zeros((J, 16), np.float32)
    for i in range(J):
        m = np.eye(4)
        m[:3, :3] = rng.normal(size=(3, 3)) * 0.7 + np.eye(3)
        m[:3, 3] = rng.uniform(-1, 1, 3)
        inv[i] = m.reshape(16)
    sk = A.SkeletonData(["b%d" % i for i in range(J)], rng.permutation(J).astype(np.int32) if seed % 5 == 0 else np.arange(J, dtype=np.int32),
                        parent, rng.uniform(-0.5, 0.5, (J, 3)).astype(np.float32), q.astype(np.float32),
                        rng.uniform(0.5, 1.5, (J, 3)).astype(np.float32), inv, True, 0, True)

    def clip(cseed, allow_dup):
        r = np.random.default_rng(cseed)
        C = int(r.integers(1, J + 6))
        total = int(r.integers(2, 14))
        pk, rk, sc = [], [], []
        for _ in range(C):
            n = [int(r.integers(1, total + 3)) if r.uniform() < 0.3 else total for _ in range(3)]
            pk.append(r.uniform(-0.2, 0.2, (n[0], 3)).astype(np.float32))
            qq = r.normal(size=(n[1], 4))
            rk.append((qq / np.linalg.norm(qq, axis=1, keepdims=True) * r.uniform(0.5, 2.0)).astype(np.float32))  # rotate() normalises
            sc.append(r.uniform(0.7, 1.3, (n[2], 3)).astype(np.float32))
        conv = r.integers(0, J, C + int(r.integers(-2, 3))).astype(np.int32) if allow_dup else r.permutation(max(J, C))[:C].astype(np.int32)
        conv = np.where(conv >= J, -1, conv).astype(np.int32)
        conv[r.uniform(size=len(conv)) < 0.1] = -1
        node_count = int(r.integers(max(1, C - 2), C + 1))
        return A.AnimationClip(total, node_count, ["c%d" % i for i in range(C)], pk, rk, sc), A.AnimSkelMap(conv)

    return sk, clip, rng


@pytest.mark.parametrize("seed", range(24))
def test_random_assets_pose_matches_oracle(oracle, seed):
    from piccolo_b200 import capi
    sk, make_clip, rng = random_assets(seed)
    ctx = capi.Context(0)
    try:
        skel = ctx.upload_skeleton(sk)
        sv = oracle.skeleton(sk)
        rc, parent, _ = oracle.build_skeleton(sv)
        assert np.array_equal(ctx.skeleton_parents(skel), parent)
        assert np.array_equal(ctx.skeleton_indices(skel), sk.index + 1)
        for allow_dup in (False, True):  # False: every bone <= 1 channel (step-ordered kernel); True: duplicates (channel-ordered kernel)
            clips = [make_clip(1000 * seed + 10 * allow_dup + k, allow_dup) for k in range(2)]
            ids = np.int32([ctx.upload_clip(skel, c, m) for c, m in clips])
            ov = [oracle.clip(c, m) for c, m in clips]
            N = 70
            ph = rng.uniform(0, 1, (N, 2)).astype(np.float32)
            ph[0], ph[1] = 0.0, 1.0
            # K = 1 on each clip
            for k in range(2):
                b = ctx.batch(skel, N, 1)
                b.keep_derived(True)
                b.set(clip_ids=np.full((N, 1), ids[k], np.int32), phase=ph[:, k:k + 1])
                b.pose()
                got, der = b.download_palette(), b.download_derived()
                ref, rder = oracle.pose_batch(sv, ov, np.full((N, 1), k, np.int32), ph[:, k:k + 1], derived=True)
                scale = np.maximum(1.0, np.abs(ref).max(axis=2, keepdims=True))
                assert (np.abs(got - ref) / scale).max() <= 2e-5, "seed %d dup %d clip %d" % (seed, allow_dup, k)
                assert np.abs(der - rder).max() <= 2e-5 * max(1.0, np.abs(rder).max())
                b.close()
            # K = 2 blend with masks
            masks = (rng.uniform(size=(2, sk.J)) < 0.8).astype(np.int32)
            mids = np.int32([ctx.upload_mask(skel, m) for m in masks])
            w = rng.uniform(0.0, 1.0, (N, 2)).astype(np.float32)
            w[3] = 0
            mp = rng.integers(0, 2, (N, 2)).astype(np.int32)
            b = ctx.batch(skel, N, 2)
            b.set(clip_ids=np.tile(ids, (N, 1)), phase=ph, weight=w, mask_ids=mids[mp])
            b.pose()
            got = b.download_palette()
            ref = oracle.pose_batch(sv, ov, np.tile(np.int32([0, 1]), (N, 1)), ph, weight=w, masks=masks, mask_ids=mp)
            scale = np.maximum(1.0, np.abs(ref).max(axis=2, keepdims=True))
            assert np.isfinite(got).all()
            assert (np.abs(got - ref) / scale).max() <= 4e-5, "seed %d dup %d blend" % (seed, allow_dup)
            b.close()
    finally:
        ctx.close()


@pytest.mark.parametrize("seed", range(12))
def test_random_meshes_skin_like_the_oracle(oracle, seed):
    """LBS fuzz: random skeleton size (2..330 bones: whole-warp, 16- and 8-instance tiles, lane-per-vertex fallback), random vertex
    count, bindings with 0..4 influences per vertex, empty slots (index 0), zero, negative and unnormalised weights, ragged
    instance counts, both output sets, whatever kernel the launcher picks."""
    from piccolo_b200 import capi
    rng = np.random.default_rng(1000 + seed)
    J = int([2, 7, 40, 64, 82, 83, 120, 165, 200, 256, 330, 58][seed])
    V = int(rng.integers(1, 6000))
    N = int([1, 5, 161, 197, 300, 163, 170, 250, 164, 180, 168, 333][seed])
    sk = A.make_synthetic_skeleton(J, branching=2 + seed % 3, seed=seed)
    cl, am = A.make_synthetic_clip(sk, 9, seed=seed + 1)
    pos = rng.uniform(-1, 1, (V, 3)).astype(np.float32)
    nrm = rng.normal(size=(V, 3)).astype(np.float32)
    tan = rng.normal(size=(V, 3)).astype(np.float32)
    idx = rng.integers(0, J + 1, (V, 4)).astype(np.int32)          # 0 = empty slot
    w = rng.uniform(0.0, 2.0, (V, 4)).astype(np.float32)
    w[rng.uniform(size=(V, 4)) < 0.3] = 0.0                         # unused influences
    w[rng.uniform(size=(V, 4)) < 0.02] *= -1.0                      # gated out by the shader's weight > 0
    if V > 3:
        w[3] = 0.0                                                  # a vertex nothing drives: position 0, normal NaN
        idx[2] = 0
    mesh = A.MeshData(pos, nrm, tan, np.zeros((V, 2), np.float32), np.arange(V, dtype=np.int32), idx, w)
    ctx = capi.Context(0)
    try:
        skel = ctx.upload_skeleton(sk)
        clip = ctx.upload_clip(skel, cl, am)
        mh = ctx.upload_mesh(mesh, J)
        b = ctx.batch(skel, N, 1)
        b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=rng.uniform(0, 1, N).astype(np.float32))
        b.pose()
        pal = b.download_palette()
        for flags in (capi.PCL_SKIN_POS_NRM_TAN, capi.PCL_SKIN_POS):
            sb = ctx.skinbuf(mh, V, N, flags)
            b.skin(mh, sb)
            gp, gnt = sb.download()
            sb.close()
            for i in sorted(set([0, N // 2, N - 1])):
                rp, rn, rt = oracle.lbs(mesh, oracle.palette_slots(pal[i]))
                radius = max(1.0, float(np.abs(rp).max()))
                assert np.abs(gp[i] - rp).max() <= 1e-4 * radius
                if gnt is not None:
                    ok = np.isfinite(rn).all(axis=1) & np.isfinite(rt).all(axis=1)
                    assert np.array_equal(np.isfinite(gnt[i]).all(axis=1), ok)
                    # normals of nearly cancelling blends are ill-conditioned: compare where the blended vector is not tiny
                    assert np.abs(gnt[i][ok, :3] - rn[ok]).max() <= 2e-3 and np.abs(gnt[i][ok, 3:] - rt[ok]).max() <= 2e-3
        b.close()
    finally:
        ctx.close()
