"""The synthetic crowds of BASELINE.json's configs (piccolo_b200/workloads.py): what bench.py runs and what the full-size GPU
tests check must be the same arrays, and a rank's shard must hold the values of the same range of the unsharded crowd
(shard == unsharded)."""
import numpy as np

from piccolo_b200 import assets as A
from piccolo_b200 import workloads as W
from piccolo_b200.sharding import shard_range


def test_shards_are_slices_of_the_unsharded_crowd():
    n = 100003
    whole3 = W.config3_inputs(0, n)
    whole4 = W.config4_inputs(0, n, total=n)
    whole2 = W.config2_inputs(0, n)
    for world in (2, 4, 8):
        parts3, parts4, parts2, covered = [], [], [], 0
        for rank in range(world):
            first, count = shard_range(n, world, rank)
            assert first == covered
            covered += count
            parts3.append(W.config3_inputs(first, count))
            parts4.append(W.config4_inputs(first, count, total=n))
            parts2.append(W.config2_inputs(first, count))
        assert covered == n
        for k in range(2):
            assert np.array_equal(np.concatenate([p[k] for p in parts3]), whole3[k])
            assert np.array_equal(np.concatenate([p[k] for p in parts4]), whole4[k])
        assert np.array_equal(np.concatenate(parts2), whole2)
    ph, w = whole3
    assert ph.dtype == np.float32 and w.dtype == np.float32 and ph.shape == (n, 2)
    assert np.all((ph >= 0) & (ph < 1)) and np.allclose(w.sum(1), 1.0)
    pick, phase = whole4
    assert pick.min() == 0 and pick.max() == 7 and np.all((phase >= 0) & (phase < 1))


def test_sample_ranges_cover_both_ends_and_do_not_overlap():
    for n_total, n_ranges, width in ((1000000, 64, 64), (320000, 64, 64), (100, 8, 64), (70, 3, 7)):
        r = W.sample_ranges(n_total, n_ranges, width)
        assert r[0][0] == 0 and (len(r) == 1 or r[-1][0] + r[-1][1] == n_total)
        idx = np.concatenate([np.arange(f, f + c) for f, c in r])
        assert len(np.unique(idx)) == len(idx) and idx.max() < n_total


def test_config5_meshes_have_four_live_influences_and_the_limb_bound_one_is_coherent():
    sk = A.make_synthetic_skeleton(64, branching=2, seed=3)
    for make in (A.make_coherent_mesh, A.make_chain_bound_mesh):
        m = make(sk, V=3000, seed=4)
        assert m.V == 3000 and np.all(m.bind_w > 0) and m.bind_idx.min() >= 1 and m.bind_idx.max() <= 64
        assert all(len(set(row)) == 4 for row in m.bind_idx)
        assert np.array_equal(m.index_buffer, np.arange(3000))  # fully expanded, as render_resource.cpp:705-712 needs
    chain = A.make_chain_bound_mesh(sk, V=3000, seed=4).bind_idx
    shared = np.mean([len(set(chain[i]) & set(chain[i - 1])) for i in range(1, 3000)])
    assert shared > 3.8  # neighbouring vertices keep (almost) all four bones
    # the player mesh the goldens were made with is untouched by the new options
    p = A.make_player_mesh(A.load_player()[0], V=600, seed=42)
    assert (p.bind_w > 0).sum(1).min() >= 1 and (p.bind_w > 0).sum(1).max() <= 4
