"""Helpers for the host C++ mirror tests: build + run piccolo_b200/host/test_host and parse its dumps."""
import json
import os
import subprocess

import numpy as np

from conftest import ROOT

DATA = "asset/objects/character/player/components/animation/data/"


def exe():
    from piccolo_b200 import build as B
    B.build_cuda()
    B.build_host()
    return B.build_host_test()


def run(args, **kw):
    return subprocess.run([exe()] + [str(a) for a in args], capture_output=True, text=True, timeout=300, **kw)


def read_dump(path):
    """-> {tag: array}; float records keep their tag, int records use tag + 'i'."""
    out, raw = {}, open(path, "rb").read()
    o = 0
    while o < len(raw):
        tag, n = np.frombuffer(raw, np.int32, 2, o)
        o += 8
        is_int = bool(tag & 0x40000000)
        tag = int(tag & 0x3FFFFFFF)
        key = "%d%s" % (tag, "i" if is_int else "")
        while key in out:  # a tag written several times: "1", "1'", "1''", ...
            key += "'"
        out[key] = np.frombuffer(raw, np.int32 if is_int else np.float32, int(n), o).copy()
        o += 4 * int(n)
    return out


def write_player_tree(td, player, sparse=False, clip_count=1, extra_clip=None):
    """Asset tree shaped like the reference's (urls relative to the root) + an object file with the
    AnimationComponent entry of player.object.json."""
    from piccolo_b200 import assets as A
    sk, cl, am, mk = player
    urls = dict(skeleton=DATA + "skeleton_data_root.skeleton.json", clip=DATA + "clip0.animation_clip.json",
                map=DATA + "anim.skeleton_map.json", mask=DATA + "anim.skeleton_mask.json")
    A.write_json(A.skeleton_to_json(sk, sparse), os.path.join(td, urls["skeleton"]))
    A.write_json(A.clip_to_json(cl, sparse), os.path.join(td, urls["clip"]))
    A.write_json(A.map_to_json(am), os.path.join(td, urls["map"]))
    A.write_json(A.mask_to_json(A.BoneBlendMask(urls["skeleton"], mk.enabled)), os.path.join(td, urls["mask"]))
    clips, lens, weights = [urls["clip"]], [0.76666998863220215], [1]
    if clip_count == 2:
        urls["clip1"] = DATA + "clip1.animation_clip.json"
        A.write_json(A.clip_to_json(extra_clip, sparse), os.path.join(td, urls["clip1"]))
        clips.append(urls["clip1"])
        lens.append(extra_clip.total_frame / 30.0)
        weights[:] = [0.25, 0.75]
    obj = {"components": [
        {"$context": {"transform": {}}, "$typeName": "TransformComponent"},
        {"$context": {"animation_res": {"blend_state": {
            "blend_anim_skel_map_path": [urls["map"]] * clip_count, "blend_clip_file_length": lens, "blend_clip_file_path": clips,
            "blend_mask_file_path": [urls["mask"], urls["mask"]], "blend_ratio": [0] * clip_count, "blend_weight": weights,
            "clip_count": clip_count}, "frame_position": 0, "skeleton_file_path": urls["skeleton"]}},
         "$typeName": "AnimationComponent"}]}
    urls["object"] = "asset/objects/character/player/player.object.json"
    with open(os.path.join(td, urls["object"]), "w") as fh:
        json.dump(obj, fh)
    return urls
