"""Host C++ mirror on the GPU: AnimationComponent / Skeleton / AnimationBatch through the reference's own call
sequence, against the reference-generated goldens and the oracle."""
import os

import numpy as np
import pytest

import hostutil as H
from piccolo_b200 import assets as A
from conftest import GOLDEN
from test_gpu_parity import assert_mats_close

pytestmark = pytest.mark.gpu


def test_animation_component_tick_sequence(player, tmp_path):
    """GObject::load + 200 x GObject::tick of the player's AnimationComponent (animation_component.cpp:8-25)."""
    g = np.load(os.path.join(GOLDEN, "tick_golden.npz"))
    urls = H.write_player_tree(str(tmp_path), player, sparse=True)
    out = tmp_path / "tick.bin"
    r = H.run(["tick", tmp_path, urls["object"], len(g["phases"]), float(g["dt"]), out])
    assert r.returncode == 0, r.stderr + r.stdout
    assert "bones=58 first=root" in r.stdout and "ctx=gpu" in r.stdout
    d = H.read_dump(out)
    assert np.array_equal(d["1"], g["phases"])  # bit-exact phase sequence
    for t, want in zip(g["mats_tick_idx"], g["mats"]):
        assert np.array_equal(d["%di" % (100 + t)], np.arange(1, 59))  # AnimationResultElement.index, bit-exact
        assert_mats_close(d[str(100 + t)].reshape(58, 16), want)
    # LevelDebugger's view: derived TRS, parents, ids
    sk = player[0]
    from oracle import pyoracle as O
    orc = O.Oracle()
    _, der = orc.pose_batch(orc.skeleton(sk), [orc.clip(player[1], player[2])], np.zeros((1, 1), np.int32), g["phases"][-1:], derived=True)
    assert np.abs(d["2"].reshape(58, 10) - der[0]).max() <= 1e-5
    rc, parent, _ = orc.build_skeleton(orc.skeleton(sk))
    assert np.array_equal(d["3i"], parent) and np.array_equal(d["4i"], sk.index)


@pytest.mark.parametrize("clip_count", [1, 2])
def test_animation_batch_crowd(player, tmp_path, clip_count):
    from oracle import pyoracle as O
    sk, cl, am, mk = player
    run_clip = A.make_variant_clip(cl, 17, seed=1234)
    urls = H.write_player_tree(str(tmp_path), player, clip_count=clip_count, extra_clip=run_clip)
    mesh = A.make_player_mesh(sk, V=3000, seed=3)
    A.write_json(A.mesh_to_json(mesh), os.path.join(str(tmp_path), "mesh.json"))
    N, ticks, dt = 777, 5, 1 / 60
    out = tmp_path / "batch.bin"
    r = H.run(["batch", tmp_path, urls["object"], N, ticks, dt, "mesh.json", out])
    assert r.returncode == 0, r.stderr + r.stdout
    d = H.read_dump(out)
    orc = O.Oracle()
    K = clip_count
    ph = np.repeat((np.arange(N, dtype=np.float32) / np.float32(N))[:, None], K, 1)
    for _ in range(ticks):  # only blend_ratio[0] advances (animation_component.cpp:19-21)
        ph[:, 0] = [orc.advance(p, np.float32(dt), np.float32(0.76666998863220215)) for p in ph[:, 0]]
    assert np.array_equal(d["1"].reshape(N, K), ph)
    clips = [orc.clip(cl, am)] + ([orc.clip(run_clip, am)] if K == 2 else [])
    w = np.tile(np.float32([1] if K == 1 else [0.25, 0.75]), (N, 1))
    ref = orc.pose_batch(orc.skeleton(sk), clips, np.tile(np.arange(K, dtype=np.int32), (N, 1)), ph, weight=w)
    assert_mats_close(d["2"].reshape(N, 58, 16), ref, tol=1e-5 if K == 1 else 2e-5)
    assert np.array_equal(d["3"].reshape(58, 16), d["2"].reshape(N, 58, 16)[-1]) and np.array_equal(d["3i"], np.arange(1, 59))
    pos = d["4"].reshape(8, mesh.V, 3)
    for i in range(8):
        rp, rn, rt = orc.lbs(mesh, orc.palette_slots(d["2"].reshape(N, 58, 16)[N - 8 + i]))
        assert np.abs(pos[i] - rp).max() <= 1e-4 * max(1.0, np.abs(rp).max())
        assert np.abs(d["5"].reshape(8, mesh.V, 6)[i][:, :3] - rn).max() <= 1e-4


@pytest.mark.parametrize("n", [1, 8, 100])
def test_components_of_a_level_pose_together(player, tmp_path, n):
    """The reference's own call pattern (object.cpp:41-50: every GObject ticks its AnimationComponent, then
    mesh_component.cpp:75 reads getResult()): n components record their ticks and the first getResult() of a frame poses all
    of them in one launch.  Every component's phases are bit-exact, its result matches the oracle, a stand-alone Skeleton
    driven with the reference's pair of calls gives the same matrices, and components joining / leaving do not disturb
    the others."""
    from oracle import pyoracle as O
    sk, cl, am, mk = player
    urls = H.write_player_tree(str(tmp_path), player, sparse=True)
    ticks, dt = 7, 1 / 60
    out = tmp_path / "multi.bin"
    r = H.run(["multi", tmp_path, urls["object"], n, ticks, dt, out])
    assert r.returncode == 0 and "multi ok" in r.stdout, r.stderr + r.stdout
    d = H.read_dump(out)
    orc = O.Oracle()
    ph = (np.arange(n, dtype=np.float32) / np.float32(n))
    for _ in range(ticks):
        ph = np.float32([orc.advance(p, np.float32(dt), np.float32(0.76666998863220215)) for p in ph])
    assert np.array_equal(d["1"], ph)  # animation_component.cpp:19-21, bit-exact
    sv, cv = orc.skeleton(sk), orc.clip(cl, am)
    ref, der = orc.pose_batch(sv, [cv], np.zeros((n, 1), np.int32), ph, derived=True)
    for i in range(n):
        assert np.array_equal(d["%di" % (100 + i)], np.arange(1, 59))
        m = d[str(100 + i)].reshape(58, 16)
        assert np.array_equal(m[:, 12:], np.tile(np.float32([0, 0, 0, 1]), (58, 1)))
        assert_mats_close(m, ref[i])
    assert np.abs(d["2"].reshape(58, 10) - der[n - 1]).max() <= 1e-5  # LevelDebugger's view of the last component
    assert np.abs(d["3"].reshape(58, 16) - d["100"].reshape(58, 16)).max() <= 2e-6  # stand-alone Skeleton, same inputs
    assert np.array_equal(d["4"], d["100"])  # component 0 after another one left and a new one joined (dt = 0 tick)
    ref0 = orc.pose_batch(sv, [cv], np.zeros((1, 1), np.int32), np.float32([0.0]))
    assert_mats_close(d["5"].reshape(58, 16), ref0[0])  # the late joiner, phase 0
