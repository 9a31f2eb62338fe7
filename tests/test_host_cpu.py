"""Host C++ mirror, CPU side: the JSON loader reproduces the reference reader's sparse-default semantics
(engine/template/allSerializer.ipp.mustache:28-33) and the classes stay silently inert without a GPU, the way the
reference stays silent on an invalid skeleton (skeleton.cpp:24-28,41-44)."""
import os

import numpy as np
import pytest

import hostutil as H
from piccolo_b200 import assets as A
from oracle import pyoracle as O


def _check_load(dump, sk, cl, am, mk):
    d = H.read_dump(dump)
    assert d["1i"].tolist() == [sk.J, int(sk.is_flat), int(sk.in_topological_order), sk.root_index, cl.total_frame, cl.node_count, cl.C]
    assert np.array_equal(d["2i"], sk.index) and np.array_equal(d["3i"], sk.parent_index)
    assert np.array_equal(d["4"].reshape(-1, 3), sk.bind_pos) and np.array_equal(d["5"].reshape(-1, 4), sk.bind_rot)
    assert np.array_equal(d["6"].reshape(-1, 3), sk.bind_scl) and np.array_equal(d["7"].reshape(-1, 16), sk.tpose)
    counts = d["8i"].reshape(-1, 3)
    assert counts[:, 0].tolist() == [k.shape[0] for k in cl.pos_keys]
    assert counts[:, 1].tolist() == [k.shape[0] for k in cl.rot_keys]
    assert np.array_equal(d["9"].reshape(-1, 3), np.concatenate(cl.pos_keys))
    assert np.array_equal(d["10"].reshape(-1, 4), np.concatenate(cl.rot_keys))
    assert np.array_equal(d["11"].reshape(-1, 3), np.concatenate(cl.scl_keys))
    assert np.array_equal(d["12i"], am.convert) and np.array_equal(d["13i"], mk.enabled)
    assert d["14i"].tolist() == [0, 0]  # missing file => default-constructed asset, no exception


@pytest.mark.parametrize("sparse", [False, True])
def test_loader_matches_fixture_bit_for_bit(player, tmp_path, sparse):
    urls = H.write_player_tree(str(tmp_path), player, sparse=sparse)
    out = tmp_path / "load.bin"
    r = H.run(["load", tmp_path, urls["skeleton"], urls["clip"], urls["map"], urls["mask"], out])
    assert r.returncode == 0, r.stderr
    assert "skeleton_path_in_mask=" + urls["skeleton"] in r.stdout
    _check_load(out, *player)


@pytest.mark.skipif(not os.path.isdir(os.path.join(O.REFERENCE_ROOT, "engine", "asset")), reason="reference tree not present")
def test_loader_on_the_bundled_reference_json(player, tmp_path):
    """The real, sparse asset files: root and 'biped' have no parent_index key, rotations are '{}', tpose lists v0..v11."""
    out = tmp_path / "load.bin"
    r = H.run(["load", os.path.join(O.REFERENCE_ROOT, "engine"), H.DATA + "skeleton_data_root.skeleton.json",
               H.DATA + "W2_CrouchWalk_Aim_F_Loop_IP.animation_clip.json", H.DATA + "anim.skeleton_map.json", H.DATA + "anim.skeleton_mask.json", out])
    assert r.returncode == 0, r.stderr
    _check_load(out, *player)


def test_binary_asset_cache_returns_exactly_what_the_json_does(player, tmp_path):
    """AnimationLoader::setCacheDir: the first load parses the JSON and writes flat images, the second is served
    from them, bit for bit; a stale (source touched), truncated or corrupted image is ignored and rewritten, never trusted."""
    import time
    urls = H.write_player_tree(str(tmp_path / "assets"), player, sparse=True)
    cache = tmp_path / "cache"
    cache.mkdir()
    args = ["load", tmp_path / "assets", urls["skeleton"], urls["clip"], urls["map"], urls["mask"]]
    dumps = []
    times = []
    for i in range(2):
        out = tmp_path / ("load%d.bin" % i)
        t0 = time.perf_counter()
        r = H.run(args + [out, cache])
        times.append(time.perf_counter() - t0)
        assert r.returncode == 0, r.stderr
        _check_load(out, *player)
        dumps.append(open(out, "rb").read())
    images = sorted(os.listdir(cache))
    assert len(images) == 4 and all(f.endswith(".pclbin") for f in images)  # skeleton, clip, map, mask; nothing for the missing file
    assert dumps[0] == dumps[1]
    # served from the cache: make the JSON unparsable but keep its size and mtime - the load must still succeed
    clip_json = tmp_path / "assets" / urls["clip"]
    st = os.stat(clip_json)
    raw = open(clip_json, "rb").read()
    open(clip_json, "wb").write(b"!" * len(raw))
    os.utime(clip_json, ns=(st.st_atime_ns, st.st_mtime_ns))
    out = tmp_path / "load_cached.bin"
    r = H.run(args + [out, cache])
    assert r.returncode == 0 and open(out, "rb").read() == dumps[0]
    # stale: restore the JSON with a new mtime -> the image no longer validates, the JSON is parsed again
    open(clip_json, "wb").write(raw)
    os.utime(clip_json, ns=(st.st_atime_ns, st.st_mtime_ns + 5_000_000_000))
    # corrupted and truncated images of the other assets
    for k, f in enumerate(images):
        p = cache / f
        b = bytearray(open(p, "rb").read())
        if k % 2 == 0:
            b[len(b) // 2] ^= 0x5a
        else:
            b = b[: len(b) // 2]
        open(p, "wb").write(bytes(b))
    out = tmp_path / "load_repaired.bin"
    r = H.run(args + [out, cache])
    assert r.returncode == 0, r.stderr
    _check_load(out, *player)
    assert open(out, "rb").read() == dumps[0]


def test_component_is_inert_without_a_gpu(player, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    urls = H.write_player_tree(str(tmp_path), player)
    out = tmp_path / "tick.bin"
    r = H.run(["tick", tmp_path, urls["object"], 3, 1 / 60, out])
    assert r.returncode == 0, r.stderr
    assert "bones=0" in r.stdout and "ctx=none" in r.stdout and "no CPU path" in r.stderr
    d = H.read_dump(out)
    assert d["100"].size == 0 and len(d["1"]) == 3  # phases still advance, results stay empty (like m_bones == nullptr)


def test_blend_weights_match_reference_and_oracle(player, oracle, tmp_path):
    """A2: AnimationManager::getBlendStateWithClipData's per-bone normalised weights (animation_system.cpp:117-149) -
    host mirror == oracle, and == the reference's own code when oracle/_ref can run here."""
    sk, cl, am, mk = player
    urls = H.write_player_tree(str(tmp_path), player)
    rng = np.random.default_rng(4)
    m0, m1 = (rng.uniform(size=sk.J) < 0.7).astype(np.int32), (rng.uniform(size=sk.J) < 0.5).astype(np.int32)
    m0[:3], m1[:3] = [1, 0, 1], [1, 1, 0]
    for name, m in (("m0", m0), ("m1", m1)):
        A.write_json(A.mask_to_json(A.BoneBlendMask(urls["skeleton"], m)), os.path.join(str(tmp_path), H.DATA + name + ".json"))
    w = (np.float32(0.3), np.float32(0.9))
    out = tmp_path / "w.bin"
    r = H.run(["weights", tmp_path, urls["skeleton"], urls["clip"], urls["map"], H.DATA + "m0.json", H.DATA + "m1.json", w[0], w[1], out])
    assert r.returncode == 0, r.stderr
    d = H.read_dump(out)
    got = np.stack([d["1"], d["1'"]])
    want = oracle.blend_weights(np.float32(w), np.stack([m0, m1]))
    both_off = (m0 == 0) & (m1 == 0)  # the reference divides 0/0 there only on paper: disabled slots are written as 0
    assert np.array_equal(got, want) and not got[:, both_off].any()
    assert d["2i"].tolist() == [2, 2, 2, cl.C]  # clips and maps are handed over by value, like the reference
    if O.RefLib.available() and os.path.isdir(O.REFERENCE_ROOT):
        ref = O.RefLib(str(tmp_path)).blend_weights(urls["skeleton"], urls["clip"], urls["map"], H.DATA + "m0.json", H.DATA + "m1.json", w[0], w[1], sk.J)
        assert np.array_equal(ref, want)


def test_offline_converter_conventions(player, tmp_path):
    """buildAnimSkelMap / addBoneBind (animation_loader.cpp:30-71, utilities.cpp:46-53): the name->index map reproduces the
    bundled anim.skeleton_map.json, bindings go to the first free of four slots as bone_index + 1."""
    sk, cl, am, mk = player
    urls = H.write_player_tree(str(tmp_path), player)
    out = tmp_path / "bind.bin"
    r = H.run(["bind", tmp_path, urls["skeleton"], urls["clip"], out])
    assert r.returncode == 0, r.stderr
    d = H.read_dump(out)
    assert np.array_equal(d["1i"][:-1], am.convert)  # the bundled map IS the name map of the bundled clip onto the skeleton
    assert d["1i"][-1] == 2147483647                 # unknown name -> INT_MAX (find_index_by_name)
    assert d["2i"].tolist() == [5, 10, 1, 58, 1, 1, 1, 1, 0, 0, 0]  # 4 slots fill in order, 5th refused, bad vertex refused
    assert d["3"].tolist() == [0.5, 0.25, 0.125, 0.0625]
