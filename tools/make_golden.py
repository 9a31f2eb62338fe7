#!/usr/bin/env python
"""Generate tests/golden/* from the reference itself.  Run in the build container (needs /root/reference):

    make -C oracle ref && python tools/make_golden.py

What it writes (all small, committed):
  player_assets.npz   the bundled player skeleton / clip / skeleton-map / mask, repacked (the JSON files
                      themselves are not copied; the GPU box has no reference tree).
  pose_golden.npz     palettes + indices (+ derived TRS) produced by the reference's own
                      Skeleton::applyAnimation / outputAnimationResult (oracle/_ref) at many phases.
  clamp_golden.npz    same, for a doctored clip (channel 3 rotation track cut to 15 keys, channel 20
                      position track cut to 9 keys) that exercises the key-count carry-over quirk
                      (skeleton.cpp:75-87).
  tick_golden.npz     200 AnimationComponent::tick steps (dt = 1/60 s) : phases + a few palettes.
  lbs_golden.npz      mesh.vert skinning restated with the reference's Matrix4x4/Vector4 operators, on a
                      2048-vertex synthetic mesh bound to the player skeleton.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200 import assets as A  # noqa: E402
from oracle import pyoracle as O  # noqa: E402

REF = O.REFERENCE_ROOT
DATA = "asset/objects/character/player/components/animation/data/"
G = os.path.join(ROOT, "tests", "golden")


def phases_config1():
    p = [k / 250.0 for k in range(250)] + [k / 22.0 for k in range(22)] + [0.0, 1.0 - 2.0 ** -24, 0.37, 0.5, 0.97]
    return np.array(p, np.float32)


def main():
    os.makedirs(G, exist_ok=True)
    eng = os.path.join(REF, "engine")
    skel = A.load_skeleton_json(os.path.join(eng, DATA, "skeleton_data_root.skeleton.json"))
    clip = A.load_clip_json(os.path.join(eng, DATA, "W2_CrouchWalk_Aim_F_Loop_IP.animation_clip.json"))
    amap = A.load_map_json(os.path.join(eng, DATA, "anim.skeleton_map.json"))
    mask = A.load_mask_json(os.path.join(eng, DATA, "anim.skeleton_mask.json"))
    A.save_player_npz(os.path.join(G, "player_assets.npz"), skel, clip, amap, mask)

    # 1. the reference on its own bundled JSON
    ref = O.RefLib(eng)
    h = ref.open(DATA + "skeleton_data_root.skeleton.json", DATA + "W2_CrouchWalk_Aim_F_Loop_IP.animation_clip.json",
                 DATA + "anim.skeleton_map.json", DATA + "anim.skeleton_mask.json")
    ph = phases_config1()
    mats, idx, der = ref.pose(h, ph, derived=True)
    k = np.arange(1, 17, dtype=np.float64)
    for p in (0.0, 0.37, 0.5, 0.97):
        i = int(np.where(ph == np.float32(p))[0][0])
        print("phase %.2f checksum %.9f" % (p, (mats[i].astype(np.float64) * k[None, :] * idx[:, None]).sum()))
    np.savez_compressed(os.path.join(G, "pose_golden.npz"), phases=ph, mats=mats, index=idx, derived=der[::16],
                        derived_phase_idx=np.arange(len(ph))[::16])

    # 2. the same assets re-emitted by OUR json writer must give the reference bit-identical results
    with tempfile.TemporaryDirectory() as td:
        for sparse in (False, True):
            sk2, cl2, am2, mk2 = A.load_player_npz(os.path.join(G, "player_assets.npz"))
            urls = O.write_player_json(td, sk2, cl2, am2, mk2, sparse=sparse, prefix="s%d_" % sparse)
            r2 = O.RefLib(td)
            h2 = r2.open(urls["skeleton"], urls["clip"], urls["map"], urls["mask"])
            m2, i2 = r2.pose(h2, ph)
            assert np.array_equal(m2, mats) and np.array_equal(i2, idx), "npz/json round trip changed the reference output"
        print("npz -> json -> reference round trip: bit-identical (dense and sparse)")

        # 3. clamp carry-over on a doctored clip
        sk2, cl2, am2, mk2 = A.load_player_npz(os.path.join(G, "player_assets.npz"))
        cl2.rot_keys[3] = cl2.rot_keys[3][:15].copy()
        cl2.pos_keys[20] = cl2.pos_keys[20][:9].copy()
        urls = O.write_player_json(td, sk2, cl2, am2, mk2, prefix="clamp_")
        r3 = O.RefLib(td)
        h3 = r3.open(urls["skeleton"], urls["clip"], urls["map"], urls["mask"])
        cph = np.array([0.0, 0.2, 0.37, 0.5, 0.62, 0.8, 0.97, 1.0 - 2.0 ** -24], np.float32)
        cm, ci = r3.pose(h3, cph)
        for p, m in zip(cph, cm):
            print("clamp phase %.2f checksum %.9f" % (p, (m.astype(np.float64) * k[None, :] * ci[:, None]).sum()))
        np.savez_compressed(os.path.join(G, "clamp_golden.npz"), phases=cph, mats=cm, index=ci,
                            cut=np.array([[3, 1, 15], [20, 0, 9]], np.int32))  # (channel, track 0=pos/1=rot/2=scl, keys kept)

    # 4. AnimationComponent::tick sequence
    tm, tp = ref.tick(h, 200, 1.0 / 60.0, 0.0)
    np.savez_compressed(os.path.join(G, "tick_golden.npz"), dt=np.float32(1.0 / 60.0), clip_len=np.float32(0.76666998863220215),
                        phases=tp, mats=tm[[0, 1, 45, 46, 199]], mats_tick_idx=np.array([0, 1, 45, 46, 199]))

    # 5. skinning
    mesh = A.make_player_mesh(skel, V=2048, seed=42)
    mesh.bind_w[5] = 0  # a vertex whose weights sum to zero: factor stays 1, B = 0, position 0 (normals NaN)
    mesh.bind_idx[6, 1] = 0  # a gated slot with non-zero weight
    i37 = int(np.where(ph == np.float32(0.37))[0][0])
    pal = np.concatenate([np.eye(4, dtype=np.float32).reshape(1, 16), mats[i37]], 0)
    op, on, ot = ref.lbs(mesh, pal)
    np.savez_compressed(os.path.join(G, "lbs_golden.npz"), pos=mesh.pos, nrm=mesh.nrm, tan=mesh.tan, idx=mesh.bind_idx,
                        w=mesh.bind_w, palette=pal, out_pos=op, out_nrm=on, out_tan=ot)
    for f in sorted(os.listdir(G)):
        print("%-22s %8d bytes" % (f, os.path.getsize(os.path.join(G, f))))


if __name__ == "__main__":
    main()
