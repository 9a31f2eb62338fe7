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
