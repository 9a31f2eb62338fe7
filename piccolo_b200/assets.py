"""Asset schema helpers for the test / bench harness (Python side).

The PRODUCT loader is C++ (piccolo_b200/host/asset_loader.cpp).  This module exists so that tests and
bench.py can (a) read the reference's JSON assets with the reference's sparse-default semantics when
/root/reference is present, (b) carry them to the GPU box as compact .npz fixtures (the reference tree
does not exist there), (c) re-emit them as JSON files for the C++ loader and for oracle/_ref, and
(d) synthesise the extra clips / meshes / skeletons that SURVEY.md section 8d prescribes because the
reference ships only one clip and no mesh.

Schema followed (reference file:line, relative to /root/reference/engine/source/runtime):
  SkeletonData / RawBone        resource/res_type/data/skeleton_data.h:9-35
  AnimationClip / Channel       resource/res_type/data/animation_clip.h:18-50
  AnimSkelMap                   resource/res_type/data/animation_skeleton_node_map.h:9-16
  BoneBlendMask                 resource/res_type/data/skeleton_mask.h:9-16
  MeshData / Vertex / Binding   resource/res_type/data/mesh_data.h:9-53
JSON reader semantics: engine/template/allSerializer.ipp.mustache:28-33 (missing key keeps the C++
default; arrays are resize()d, i.e. value-initialised), numbers parsed as double then cast to float
(core/meta/serializer/serializer.cpp:47-52).
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

F32 = np.float32
INT_MAX = 2147483647


# --------------------------------------------------------------------------------------------
# schema
# --------------------------------------------------------------------------------------------
@dataclass
class SkeletonData:
    names: List[str]
    index: np.ndarray  # int32 [J]
    parent_index: np.ndarray  # int32 [J]
    bind_pos: np.ndarray  # f32 [J,3]
    bind_rot: np.ndarray  # f32 [J,4]  (w,x,y,z) raw, NOT normalised
    bind_scl: np.ndarray  # f32 [J,3]
    tpose: np.ndarray  # f32 [J,16] row-major inverse bind ("tpose_matrix")
    is_flat: bool = False
    root_index: int = 0
    in_topological_order: bool = False

    @property
    def J(self) -> int:
        return len(self.names)


@dataclass
class AnimationClip:
    total_frame: int
    node_count: int
    names: List[str]
    pos_keys: List[np.ndarray]  # per channel f32 [n,3]
    rot_keys: List[np.ndarray]  # per channel f32 [n,4] (w,x,y,z)
    scl_keys: List[np.ndarray]  # per channel f32 [n,3]
    node_map: List[str] = field(default_factory=list)  # AnimationAsset.node_map.convert
    skeleton_file_path: str = ""

    @property
    def C(self) -> int:
        return len(self.names)


@dataclass
class AnimSkelMap:
    convert: np.ndarray  # int32 [C]


@dataclass
class BoneBlendMask:
    skeleton_file_path: str
    enabled: np.ndarray  # int32 [J]


@dataclass
class MeshData:
    pos: np.ndarray  # f32 [V,3]
    nrm: np.ndarray  # f32 [V,3]
    tan: np.ndarray  # f32 [V,3]
    uv: np.ndarray  # f32 [V,2]
    index_buffer: np.ndarray  # int32 [I]
    bind_idx: np.ndarray  # int32 [V,4]   bone index + 1, 0 = empty slot (animation_loader.cpp:40-71)
    bind_w: np.ndarray  # f32 [V,4]   raw weights (renormalised at upload, render_resource.cpp:714-725)

    @property
    def V(self) -> int:
        return self.pos.shape[0]


# --------------------------------------------------------------------------------------------
# JSON -> schema (reference reader semantics)
# --------------------------------------------------------------------------------------------
def _f(d: dict, k: str, default: float) -> np.float32:
    v = d.get(k)
    return F32(default if v is None else v)


def _vec3(d: Optional[dict], default=(0.0, 0.0, 0.0)):
    d = d or {}
    return [_f(d, "x", default[0]), _f(d, "y", default[1]), _f(d, "z", default[2])]


def _quat(d: Optional[dict]):
    d = d or {}
    return [_f(d, "w", 1.0), _f(d, "x", 0.0), _f(d, "y", 0.0), _f(d, "z", 0.0)]


_MAT_DEFAULT = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]


def load_skeleton_json(path: str) -> SkeletonData:
    with open(path) as fh:
        j = json.load(fh)
    bones = j.get("bones_map") or []
    J = len(bones)
    names, index, parent = [], np.zeros(J, np.int32), np.zeros(J, np.int32)
    pos, rot, scl = np.zeros((J, 3), F32), np.zeros((J, 4), F32), np.zeros((J, 3), F32)
    tp = np.zeros((J, 16), F32)
    for i, b in enumerate(bones):
        names.append(b.get("name", ""))
        index[i] = int(b.get("index", 0))  # value-initialised RawBone => 0
        parent[i] = int(b.get("parent_index", 0))
        bp = b.get("binding_pose") or {}
        pos[i] = _vec3(bp.get("position"))
        scl[i] = _vec3(bp.get("scale"), (1.0, 1.0, 1.0))
        rot[i] = _quat(bp.get("rotation"))
        m = b.get("tpose_matrix") or {}
        tp[i] = [_f(m, "v%d" % k, _MAT_DEFAULT[k]) for k in range(16)]
    return SkeletonData(names, index, parent, pos, rot, scl, tp, bool(j.get("is_flat", False)),
                        int(j.get("root_index", 0)), bool(j.get("in_topological_order", False)))


def load_clip_json(path: str) -> AnimationClip:
    with open(path) as fh:
        j = json.load(fh)
    c = j.get("clip_data") or {}
    chans = c.get("node_channels") or []
    names, pk, rk, sk = [], [], [], []
    for ch in chans:
        names.append(ch.get("name", ""))
        pk.append(np.array([_vec3(k) for k in (ch.get("position_keys") or [])], F32).reshape(-1, 3))
        rk.append(np.array([_quat(k) for k in (ch.get("rotation_keys") or [])], F32).reshape(-1, 4))
        sk.append(np.array([_vec3(k) for k in (ch.get("scaling_keys") or [])], F32).reshape(-1, 3))
    return AnimationClip(int(c.get("total_frame", 0)), int(c.get("node_count", 0)), names, pk, rk, sk,
                         list((j.get("node_map") or {}).get("convert") or []), j.get("skeleton_file_path", ""))


def load_map_json(path: str) -> AnimSkelMap:
    with open(path) as fh:
        j = json.load(fh)
    return AnimSkelMap(np.array([int(x) for x in (j.get("convert") or [])], np.int32))


def load_mask_json(path: str) -> BoneBlendMask:
    with open(path) as fh:
        j = json.load(fh)
    return BoneBlendMask(j.get("skeleton_file_path", ""), np.array([int(x) for x in (j.get("enabled") or [])], np.int32))


def load_mesh_json(path: str) -> MeshData:
    with open(path) as fh:
        j = json.load(fh)
    vb = j.get("vertex_buffer") or []
    V = len(vb)
    a = np.zeros((V, 11), F32)
    keys = ["px", "py", "pz", "nx", "ny", "nz", "tx", "ty", "tz", "u", "v"]
    for i, v in enumerate(vb):
        a[i] = [F32(v.get(k, 0.0)) for k in keys]
    bind = j.get("bind") or []
    bi, bw = np.zeros((len(bind), 4), np.int32), np.zeros((len(bind), 4), F32)
    for i, b in enumerate(bind):
        bi[i] = [int(b.get("index%d" % k, 0)) for k in range(4)]
        bw[i] = [F32(b.get("weight%d" % k, 0.0)) for k in range(4)]
    return MeshData(a[:, 0:3].copy(), a[:, 3:6].copy(), a[:, 6:9].copy(), a[:, 9:11].copy(),
                    np.array([int(x) for x in (j.get("index_buffer") or [])], np.int32), bi, bw)


# --------------------------------------------------------------------------------------------
# schema -> JSON (what the reference's Serializer::write would emit; `sparse` drops default values
# the way the bundled assets do, to exercise the missing-key path of the loaders)
# --------------------------------------------------------------------------------------------
def _r(x) -> float:
    # shortest decimal string that round-trips the float32 exactly (parsed back as double -> float)
    return float(np.format_float_positional(F32(x), unique=True, trim="0"))


def _obj(keys, vals, defaults, sparse):
    return {k: _r(v) for k, v, d in zip(keys, vals, defaults) if not (sparse and F32(v) == F32(d))}


def skeleton_to_json(s: SkeletonData, sparse: bool = False) -> dict:
    bones = []
    for i in range(s.J):
        b = {"name": s.names[i], "index": int(s.index[i])}
        b["binding_pose"] = {
            "position": _obj("xyz", s.bind_pos[i], (0, 0, 0), sparse),
            "scale": _obj("xyz", s.bind_scl[i], (1, 1, 1), sparse),
            "rotation": _obj("wxyz", s.bind_rot[i], (1, 0, 0, 0), sparse),
        }
        b["tpose_matrix"] = _obj(["v%d" % k for k in range(16)], s.tpose[i], _MAT_DEFAULT, sparse)
        if not (sparse and int(s.parent_index[i]) == 0):
            b["parent_index"] = int(s.parent_index[i])
        bones.append(b)
    return {"bones_map": bones, "is_flat": bool(s.is_flat), "root_index": int(s.root_index),
            "in_topological_order": bool(s.in_topological_order)}


def clip_to_json(c: AnimationClip, sparse: bool = False) -> dict:
    chans = []
    for i in range(c.C):
        chans.append({
            "name": c.names[i],
            "position_keys": [_obj("xyz", k, (0, 0, 0), sparse) for k in c.pos_keys[i]],
            "rotation_keys": [_obj("wxyz", k, (1, 0, 0, 0), sparse) for k in c.rot_keys[i]],
            "scaling_keys": [_obj("xyz", k, (0, 0, 0), sparse) for k in c.scl_keys[i]],
        })
    return {"node_map": {"convert": list(c.node_map)},
            "clip_data": {"total_frame": int(c.total_frame), "node_count": int(c.node_count), "node_channels": chans},
            "skeleton_file_path": c.skeleton_file_path}


def map_to_json(m: AnimSkelMap) -> dict:
    return {"convert": [int(x) for x in m.convert]}


def mask_to_json(m: BoneBlendMask) -> dict:
    return {"skeleton_file_path": m.skeleton_file_path, "enabled": [int(x) for x in m.enabled]}


def mesh_to_json(m: MeshData) -> dict:
    keys = ["px", "py", "pz", "nx", "ny", "nz", "tx", "ty", "tz", "u", "v"]
    a = np.concatenate([m.pos, m.nrm, m.tan, m.uv], axis=1)
    vb = [{k: _r(x) for k, x in zip(keys, row)} for row in a]
    bind = []
    for i in range(m.bind_idx.shape[0]):
        d = {"index%d" % k: int(m.bind_idx[i, k]) for k in range(4)}
        d.update({"weight%d" % k: _r(m.bind_w[i, k]) for k in range(4)})
        bind.append(d)
    return {"vertex_buffer": vb, "index_buffer": [int(x) for x in m.index_buffer], "bind": bind}


def write_json(obj: dict, path: str) -> str:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as fh:
        json.dump(obj, fh)
    return path


# --------------------------------------------------------------------------------------------
# .npz fixtures (compact carriers of the bundled player assets; made by tools/make_golden.py)
# --------------------------------------------------------------------------------------------
def _ragged_pack(lst, width):
    counts = np.array([a.shape[0] for a in lst], np.int32)
    flat = np.concatenate([a.reshape(-1, width) for a in lst], axis=0).astype(F32) if lst else np.zeros((0, width), F32)
    return counts, flat


def _ragged_unpack(counts, flat):
    out, o = [], 0
    for n in counts:
        out.append(flat[o:o + int(n)].copy())
        o += int(n)
    return out


def save_player_npz(path, skel: SkeletonData, clip: AnimationClip, amap: AnimSkelMap, mask: BoneBlendMask):
    pc, pf = _ragged_pack(clip.pos_keys, 3)
    rc, rf = _ragged_pack(clip.rot_keys, 4)
    sc, sf = _ragged_pack(clip.scl_keys, 3)
    np.savez_compressed(
        path, skel_names=np.array(skel.names), skel_index=skel.index, skel_parent=skel.parent_index,
        skel_pos=skel.bind_pos, skel_rot=skel.bind_rot, skel_scl=skel.bind_scl, skel_tpose=skel.tpose,
        skel_flags=np.array([skel.is_flat, skel.in_topological_order, skel.root_index], np.int32),
        clip_meta=np.array([clip.total_frame, clip.node_count], np.int32), clip_names=np.array(clip.names),
        clip_node_map=np.array(clip.node_map), clip_skel_path=np.array(clip.skeleton_file_path),
        clip_pc=pc, clip_pf=pf, clip_rc=rc, clip_rf=rf, clip_sc=sc, clip_sf=sf,
        map_convert=amap.convert, mask_enabled=mask.enabled, mask_skel_path=np.array(mask.skeleton_file_path))


def load_player_npz(path):
    z = np.load(path, allow_pickle=False)
    fl = z["skel_flags"]
    skel = SkeletonData([str(x) for x in z["skel_names"]], z["skel_index"].astype(np.int32), z["skel_parent"].astype(np.int32),
                        z["skel_pos"].astype(F32), z["skel_rot"].astype(F32), z["skel_scl"].astype(F32),
                        z["skel_tpose"].astype(F32), bool(fl[0]), int(fl[2]), bool(fl[1]))
    clip = AnimationClip(int(z["clip_meta"][0]), int(z["clip_meta"][1]), [str(x) for x in z["clip_names"]],
                         _ragged_unpack(z["clip_pc"], z["clip_pf"]), _ragged_unpack(z["clip_rc"], z["clip_rf"]),
                         _ragged_unpack(z["clip_sc"], z["clip_sf"]), [str(x) for x in z["clip_node_map"]],
                         str(z["clip_skel_path"]))
    amap = AnimSkelMap(z["map_convert"].astype(np.int32))
    mask = BoneBlendMask(str(z["mask_skel_path"]), z["mask_enabled"].astype(np.int32))
    return skel, clip, amap, mask


GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def load_player():
    """The bundled player skeleton + clip + map + mask, from the committed fixture."""
    return load_player_npz(os.path.join(GOLDEN_DIR, "player_assets.npz"))


# --------------------------------------------------------------------------------------------
# float32 helpers shared by the synthetic generators
# --------------------------------------------------------------------------------------------
def quat_mul(a, b):
    """(w,x,y,z) Hamilton product, float64 (generator-side only)."""
    w1, x1, y1, z1 = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    w2, x2, y2, z2 = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 + y1 * w2 + z1 * x2 - x1 * z2, w1 * z2 + z1 * w2 + x1 * y2 - y1 * x2], axis=-1)


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def bind_model_transforms(skel: SkeletonData):
    """float64 model-space bind TRS per bone (same TRS composition rule as Node::updateDerivedTransform)."""
    J = skel.J
    P, Q, S = np.zeros((J, 3)), np.zeros((J, 4)), np.zeros((J, 3))
    for i in range(J):
        q = skel.bind_rot[i].astype(np.float64)
        q = q / np.linalg.norm(q)
        p, s = skel.bind_pos[i].astype(np.float64), skel.bind_scl[i].astype(np.float64)
        par = int(skel.parent_index[i])
        if par == INT_MAX or par >= i:
            P[i], Q[i], S[i] = p, q, s
        else:
            Q[i] = quat_mul(Q[par], q)
            Q[i] /= np.linalg.norm(Q[i])
            S[i] = S[par] * s
            P[i] = quat_to_mat(Q[par]) @ (S[par] * p) + P[par]
    return P, Q, S


# --------------------------------------------------------------------------------------------
# synthetic assets (the reference ships one clip and no mesh)
# --------------------------------------------------------------------------------------------
def make_variant_clip(base: AnimationClip, frames: int, seed: int, max_deg: float = 10.0) -> AnimationClip:
    """Time-warp `base` to `frames` frames and perturb rotations by seeded small-angle quaternions."""
    rng = np.random.default_rng(seed)
    pk, rk, sk = [], [], []
    for c in range(base.C):
        def resample(keys, width):
            n = keys.shape[0]
            t = np.linspace(0.0, n - 1, frames)
            lo = np.floor(t).astype(int)
            hi = np.minimum(lo + 1, n - 1)
            r = (t - lo)[:, None]
            return (keys[lo].astype(np.float64) * (1 - r) + keys[hi].astype(np.float64) * r)
        p = resample(base.pos_keys[c], 3)
        s = resample(base.scl_keys[c], 3)
        q = base.rot_keys[c].astype(np.float64)
        for k in range(1, q.shape[0]):  # make consecutive keys sign-consistent before resampling
            if np.dot(q[k], q[k - 1]) < 0:
                q[k:] *= -1
        q = resample(q.astype(F32), 4)
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        ang = np.deg2rad(max_deg) * rng.uniform(-1, 1) * np.sin(np.linspace(0, 2 * np.pi, frames))
        dq = np.concatenate([np.cos(ang / 2)[:, None], np.sin(ang / 2)[:, None] * axis[None, :]], axis=1)
        q = quat_mul(q, dq)
        pk.append(p.astype(F32)), rk.append(q.astype(F32)), sk.append(s.astype(F32))
    return AnimationClip(frames, base.node_count, list(base.names), pk, rk, sk, list(base.node_map), base.skeleton_file_path)


def make_player_mesh(skel: SkeletonData, V: int = 16380, seed: int = 42, full4: bool = False, chain: bool = False) -> MeshData:
    """Synthetic stand-in for the missing robot.mesh_bind.json: a tessellated tube
    (capsule side) around every bone, emitted as FULLY EXPANDED triangles in strip order - three vertices per
    triangle, index_buffer[i] == i, which is the only case in which the reference's per-index binding upload is
    self-consistent (render_resource.cpp:705-712).  Every vertex is bound to the bones among its 4 nearest whose
    distance is within 1.6x the nearest one (1-4 influences, inverse-distance weights), so - like any real
    character mesh - neighbouring vertices share most of their bones.  Indices are stored +1, 0 = empty slot
    (animation_loader.cpp:40-71)."""
    rng = np.random.default_rng(seed)
    P, _, _ = bind_model_transforms(skel)
    J = skel.J
    n_tri = max(1, V // 3)
    per = np.full(J, n_tri // J)
    per[: n_tri - per.sum()] += 1
    par = np.array([p if (p != INT_MAX and p < i) else i for i, p in enumerate(skel.parent_index)])
    pos_l, nrm_l, tan_l = [], [], []
    for b in range(J):
        nt = int(per[b])
        if nt == 0:
            continue
        a0, a1 = P[par[b]], P[b]
        axis = a1 - a0
        L = np.linalg.norm(axis)
        axis = axis / L if L > 1e-6 else np.array([0.0, 0.0, 1.0])
        L = max(L, 0.05)
        ref = np.array([1.0, 0, 0]) if abs(axis[0]) < 0.9 else np.array([0, 1.0, 0])
        e1 = np.cross(axis, ref)
        e1 /= np.linalg.norm(e1)
        e2 = np.cross(axis, e1)
        radius = 0.03 + 0.02 * rng.uniform()
        seg = 8
        rings = max(1, int(np.ceil(nt / (2 * seg))))

        def surf(ring, s):
            th = 2 * np.pi * (s % seg) / seg
            d = np.cos(th) * e1 + np.sin(th) * e2
            return a0 + axis * (L * ring / rings) + radius * d, d, axis

        tris = []
        for r in range(rings):
            for s_ in range(seg):
                q00, q01, q10, q11 = surf(r, s_), surf(r, s_ + 1), surf(r + 1, s_), surf(r + 1, s_ + 1)
                tris.append((q00, q01, q10))
                tris.append((q01, q11, q10))
        for t in tris[:nt]:
            for pnt, d, ax in t:
                pos_l.append(pnt), nrm_l.append(d), tan_l.append(ax)
    pos, nrm, tan = np.array(pos_l), np.array(nrm_l), np.array(tan_l)
    # V may not be a multiple of 3: top up with copies of the last vertex / trim
    if pos.shape[0] < V:
        k = V - pos.shape[0]
        pos, nrm, tan = (np.concatenate([a, np.repeat(a[-1:], k, 0)], 0) for a in (pos, nrm, tan))
    pos, nrm, tan = pos[:V], nrm[:V], tan[:V]
    nearest, dn = np.zeros((V, 4), np.int64), np.zeros((V, 4))
    for v0 in range(0, V, 8192):  # [V,J] distances in slabs (a 50 k x 256 mesh would need 300 MB at once)
        d = np.linalg.norm(pos[v0:v0 + 8192, None, :] - P[None, :, :], axis=2)
        nearest[v0:v0 + 8192] = np.argsort(d, axis=1, kind="stable")[:, :4]
        dn[v0:v0 + 8192] = np.take_along_axis(d, nearest[v0:v0 + 8192], 1)
    on = (dn <= 1.6 * dn[:, :1] + 1e-9) | full4
    if chain:
        # bind every vertex to the bones AROUND the tube it lies on - that bone, its parent, its grandparent and its first
        # child - which is how skinning weights are painted on a real character (influences follow the limb, not a sphere
        # in space); inverse-distance weights, all four non-zero
        first_child = np.full(J, -1)
        for b in range(J - 1, 0, -1):
            if par[b] != b:
                first_child[par[b]] = b
        owner = np.repeat(np.arange(J), per * 3)[:V]
        owner = np.concatenate([owner, np.full(V - len(owner), owner[-1] if len(owner) else 0)])
        cand = np.stack([owner, par[owner], par[par[owner]], np.where(first_child[owner] >= 0, first_child[owner], par[par[par[owner]]])], 1)
        for v_, row in enumerate(cand):  # four DISTINCT bones: near the root the chain folds onto itself
            seen = []
            for b in row:
                while int(b) in seen:
                    b = (int(b) + 1) % J
                seen.append(int(b))
            cand[v_] = seen
        nearest = cand
        dn = np.linalg.norm(pos[:, None, :] - P[nearest], axis=2)
        on = np.ones_like(dn, bool)
    idx = np.where(on, nearest + 1, 0).astype(np.int32)
    w = np.where(on, 1.0 / (dn + 1e-3), 0.0).astype(F32)
    uv = rng.uniform(0, 1, size=(V, 2))
    return MeshData(pos.astype(F32), nrm.astype(F32), tan.astype(F32), uv.astype(F32), np.arange(V, dtype=np.int32), idx, w)


def make_coherent_mesh(skel: SkeletonData, V: int = 49998, seed: int = 11) -> MeshData:
    """A character-like mesh for any skeleton: make_player_mesh's tubes around the bones, every vertex bound to its FOUR
    nearest bones with non-zero inverse-distance weights (config 5 asks for 4 non-zero weights per vertex)."""
    return make_player_mesh(skel, V=V, seed=seed, full4=True)


def make_chain_bound_mesh(skel: SkeletonData, V: int = 49998, seed: int = 11) -> MeshData:
    """make_player_mesh's tubes with every vertex bound to the bones around its tube (bone, parent, grandparent, first child):
    four non-zero weights that change only where the mesh moves on to another bone, like a painted character."""
    return make_player_mesh(skel, V=V, seed=seed, chain=True)


def make_synthetic_skeleton(J: int = 256, branching: int = 3, seed: int = 5) -> SkeletonData:
    """Balanced tree, topological order, random unit bind quaternions, offsets in [-0.2,0.2]^3, unit scale,
    inverse bind = numerically inverted bind model matrix (config 5)."""
    rng = np.random.default_rng(seed)
    parent = np.array([0] + [(i - 1) // branching for i in range(1, J)], np.int32)
    parent[0] = INT_MAX
    q = rng.normal(size=(J, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    pos = rng.uniform(-0.2, 0.2, size=(J, 3))
    pos[0] = 0
    skel = SkeletonData(["joint_%03d" % i for i in range(J)], np.arange(J, dtype=np.int32), parent,
                        pos.astype(F32), q.astype(F32), np.ones((J, 3), F32), np.zeros((J, 16), F32), True, 0, True)
    P, Q, S = bind_model_transforms(skel)
    for i in range(J):
        M = np.eye(4)
        M[:3, :3] = quat_to_mat(Q[i]) * S[i][None, :]
        M[:3, 3] = P[i]
        inv = np.linalg.inv(M)
        inv[3] = [0, 0, 0, 1]
        skel.tpose[i] = inv.reshape(16).astype(F32)
    return skel


def make_synthetic_clip(skel: SkeletonData, frames: int = 32, seed: int = 6):
    """One channel per bone (identity map), looping small rotations/translations, non-unit scale keys."""
    rng = np.random.default_rng(seed)
    J = skel.J
    pk, rk, sk = [], [], []
    tt = np.linspace(0, 2 * np.pi, frames)
    for _ in range(J):
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        ang = rng.uniform(0.1, 0.6) * np.sin(tt + rng.uniform(0, 6.28))
        rk.append(np.concatenate([np.cos(ang / 2)[:, None], np.sin(ang / 2)[:, None] * axis[None, :]], axis=1).astype(F32))
        pk.append((rng.uniform(-0.02, 0.02, size=(1, 3)) * np.sin(tt)[:, None]).astype(F32))
        sk.append((1.0 + 0.05 * rng.uniform(-1, 1, size=(1, 3)) * np.cos(tt)[:, None]).astype(F32))
    clip = AnimationClip(frames, J, list(skel.names), pk, rk, sk, list(skel.names), "")
    return clip, AnimSkelMap(np.arange(J, dtype=np.int32))


def make_synthetic_mesh(skel: SkeletonData, V: int = 49998, seed: int = 9) -> MeshData:
    """Config 5 stress mesh: 4 non-zero weights per vertex."""
    rng = np.random.default_rng(seed)
    P, _, _ = bind_model_transforms(skel)
    J = skel.J
    owner = rng.integers(0, J, size=V)
    u = rng.normal(size=(V, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    pos = P[owner] + u * 0.05
    tan = np.cross(u, np.roll(u, 1, axis=1))
    tan /= np.maximum(np.linalg.norm(tan, axis=1, keepdims=True), 1e-9)
    idx = np.zeros((V, 4), np.int32)
    idx[:, 0] = owner + 1
    for s in range(1, 4):
        idx[:, s] = (owner + rng.integers(1, 8, size=V)) % J + 1
    w = rng.uniform(0.1, 1.0, size=(V, 4)).astype(F32)
    return MeshData(pos.astype(F32), u.astype(F32), tan.astype(F32), rng.uniform(0, 1, size=(V, 2)).astype(F32),
                    np.arange(V, dtype=np.int32), idx, w)
