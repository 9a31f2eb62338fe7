"""ctypes bindings for the two CPU checkers.  TEST INFRASTRUCTURE ONLY.

  Oracle()   -> oracle/liboracle.so           plain-C restatement (oracle/oracle.c)
  RefLib()   -> oracle/_ref/libpiccolo_ref.so the unmodified reference Skeleton code (oracle/Makefile)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpiccolo_ref.so")
REFERENCE_ROOT = os.environ.get("PICCOLO_REF", "/root/reference")

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)


def _f(a):
    return a.ctypes.data_as(_fp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def build(ref: bool = True) -> None:
    """Compile the checkers (gcc/g++ only).  oracle/_ref needs the reference tree and is skipped without it."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref and os.path.isdir(os.path.join(REFERENCE_ROOT, "engine", "source", "runtime")):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref", "PICCOLO_REF=" + REFERENCE_ROOT])


class _Skel(C.Structure):
    _fields_ = [("J", C.c_int32), ("index", _ip), ("parent", _ip), ("bind_pos", _fp), ("bind_rot", _fp),
                ("bind_scl", _fp), ("inv_bind", _fp), ("is_flat", C.c_int32), ("in_topo", C.c_int32)]


class _Clip(C.Structure):
    _fields_ = [("total_frame", C.c_int32), ("node_count", C.c_int32), ("C", C.c_int32), ("convert_len", C.c_int32),
                ("key_counts", _ip), ("key_offsets", _ip), ("pos", _fp), ("rot", _fp), ("scl", _fp), ("convert", _ip)]


class Oracle:
    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        self.lib.orc_advance.restype = C.c_float
        self.lib.orc_advance.argtypes = [C.c_float, C.c_float, C.c_float]
        self._keep = []

    # -- views ---------------------------------------------------------------------------
    def skeleton(self, s):
        a = dict(index=np.ascontiguousarray(s.index, np.int32), parent=np.ascontiguousarray(s.parent_index, np.int32),
                 pos=np.ascontiguousarray(s.bind_pos, np.float32), rot=np.ascontiguousarray(s.bind_rot, np.float32),
                 scl=np.ascontiguousarray(s.bind_scl, np.float32), inv=np.ascontiguousarray(s.tpose, np.float32))
        v = _Skel(s.J, _i(a["index"]), _i(a["parent"]), _f(a["pos"]), _f(a["rot"]), _f(a["scl"]), _f(a["inv"]),
                  int(s.is_flat), int(s.in_topological_order))
        self._keep.append(a)
        return v

    def clip(self, c, amap):
        counts = np.zeros((c.C, 3), np.int32)
        offs = np.zeros((c.C, 3), np.int32)
        o = [0, 0, 0]
        for n in range(c.C):
            for t, keys in enumerate((c.pos_keys[n], c.rot_keys[n], c.scl_keys[n])):
                counts[n, t] = keys.shape[0]
                offs[n, t] = o[t]
                o[t] += keys.shape[0]
        cat = lambda lst, w: (np.ascontiguousarray(np.concatenate([k.reshape(-1, w) for k in lst], 0), np.float32)
                              if lst else np.zeros((0, w), np.float32))
        a = dict(counts=counts, offs=offs, pos=cat(c.pos_keys, 3), rot=cat(c.rot_keys, 4), scl=cat(c.scl_keys, 3),
                 conv=np.ascontiguousarray(amap.convert, np.int32))
        v = _Clip(c.total_frame, c.node_count, c.C, len(a["conv"]), _i(a["counts"]), _i(a["offs"]), _f(a["pos"]),
                  _f(a["rot"]), _f(a["scl"]), _i(a["conv"]))
        self._keep.append(a)
        return v

    # -- calls -----------------------------------------------------------------------------
    def build_skeleton(self, sv):
        parent = np.zeros(sv.J, np.int32)
        rot = np.zeros((sv.J, 4), np.float32)
        rc = self.lib.orc_build(C.byref(sv), _i(parent), _f(rot))
        return rc, parent, rot

    def advance(self, ratio, dt, clip_len):
        return np.float32(self.lib.orc_advance(np.float32(ratio), np.float32(dt), np.float32(clip_len)))

    def blend_weights(self, weight, enabled):
        K, J = enabled.shape
        out = np.zeros((K, J), np.float32)
        w = np.ascontiguousarray(weight, np.float32)
        e = np.ascontiguousarray(enabled, np.int32)
        self.lib.orc_blend_weights(K, J, _f(w), _i(e), _f(out))
        return out

    def pose_batch(self, sv, clip_views, clip_ids, phase, weight=None, masks=None, mask_ids=None, derived=False,
                   n_threads=None):
        """clip_ids/phase/weight: [N,K].  Returns mats [N,J,16] (+ derived [N,J,10])."""
        clip_ids = np.ascontiguousarray(clip_ids, np.int32).reshape(len(clip_ids), -1)
        N, K = clip_ids.shape
        phase = np.ascontiguousarray(phase, np.float32).reshape(N, K)
        if weight is None:
            weight = np.ones((N, K), np.float32)
        weight = np.ascontiguousarray(weight, np.float32).reshape(N, K)
        arr = (C.POINTER(_Clip) * len(clip_views))(*[C.pointer(c) for c in clip_views])
        mats = np.zeros((N, sv.J, 16), np.float32)
        der = np.zeros((N, sv.J, 10), np.float32) if derived else None
        m = np.ascontiguousarray(masks, np.int32) if masks is not None else None
        mi = np.ascontiguousarray(mask_ids, np.int32).reshape(N, K) if mask_ids is not None else None
        nt = n_threads or min(os.cpu_count() or 1, max(1, N // 64))
        rc = self.lib.orc_pose_batch(C.byref(sv), len(clip_views), arr, 0 if m is None else m.shape[0], _i(m), N, K,
                                     _i(clip_ids), _f(phase), _f(weight), _i(mi), _f(mats), _f(der), nt)
        if rc != 0:
            raise RuntimeError("orc_pose_batch: skeleton rejected (not flat / not topological), rc=%d" % rc)
        return (mats, der) if derived else mats

    def indices(self, sv):
        return np.ctypeslib.as_array(sv.index, shape=(sv.J,)).astype(np.int32) + 1

    def normalise_weights(self, w_raw):
        w_raw = np.ascontiguousarray(w_raw, np.float32)
        out = np.zeros_like(w_raw)
        self.lib.orc_normalise_weights(w_raw.shape[0], _f(w_raw), _f(out))
        return out

    def palette_slots(self, mats):
        """[J,16] -> [J+1,16] with the identity in slot 0 (mesh_component.cpp:72-79)."""
        J = mats.shape[0]
        out = np.zeros((J + 1, 16), np.float32)
        m = np.ascontiguousarray(mats, np.float32)
        self.lib.orc_palette_slots(J, _f(m), _f(out))
        return out

    def lbs(self, mesh, palette16, normals=True):
        """mesh: assets.MeshData (raw weights).  palette16: [J+1,16].  Returns pos (,nrm,tan) [V,3]."""
        w = self.normalise_weights(mesh.bind_w)
        V = mesh.V
        pos = np.ascontiguousarray(mesh.pos, np.float32)
        nrm = np.ascontiguousarray(mesh.nrm, np.float32)
        tan = np.ascontiguousarray(mesh.tan, np.float32)
        idx = np.ascontiguousarray(mesh.bind_idx, np.int32)
        pal = np.ascontiguousarray(palette16, np.float32)
        op = np.zeros((V, 3), np.float32)
        on = np.zeros((V, 3), np.float32) if normals else None
        ot = np.zeros((V, 3), np.float32) if normals else None
        self.lib.orc_lbs(V, _f(pos), _f(nrm), _f(tan), _i(idx), _f(w), _f(pal), _f(op), _f(on), _f(ot))
        return (op, on, ot) if normals else op


class RefLib:
    """The reference's own Skeleton / AnimationManager code.  Needs asset JSON files on disk."""

    def __init__(self, asset_root: str):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        self.lib = L = C.CDLL(REF_SO)
        self.root = asset_root
        L.ref_set_asset_root(asset_root.encode())
        L.ref_open.argtypes = [C.c_char_p] * 4 + [C.c_float]
        L.ref_pose.argtypes = [C.c_int, C.c_int, _fp, _fp, _ip, _fp]
        L.ref_tick.argtypes = [C.c_int, C.c_int, C.c_float, _fp, _fp, _fp]
        L.ref_bench_pose.restype = C.c_double
        L.ref_bench_pose.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_double)]
        L.ref_bench_lbs.restype = C.c_double
        L.ref_bench_lbs.argtypes = [C.c_int, _fp, _fp, _fp, _ip, _fp, C.c_int, _fp, C.c_int, C.c_int, C.POINTER(C.c_double)]

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def open(self, skeleton_url, clip_url, map_url, mask_url, clip_len=0.76666998863220215):
        self.lib.ref_set_asset_root(self.root.encode())
        return self.lib.ref_open(skeleton_url.encode(), clip_url.encode(), map_url.encode(), mask_url.encode(),
                                 C.c_float(clip_len))

    def bone_count(self, h):
        return self.lib.ref_bone_count(h)

    def pose(self, h, phases, derived=False):
        ph = np.ascontiguousarray(phases, np.float32)
        J = self.bone_count(h)
        mats = np.zeros((len(ph), J, 16), np.float32)
        idx = np.zeros(J, np.int32)
        der = np.zeros((len(ph), J, 10), np.float32) if derived else None
        rc = self.lib.ref_pose(h, len(ph), _f(ph), _f(mats), _i(idx), _f(der))
        if rc != 0:
            raise RuntimeError("ref_pose failed")
        return (mats, idx, der) if derived else (mats, idx)

    def tick(self, h, n_ticks, dt, phase0):
        J = self.bone_count(h)
        ph = np.array([phase0], np.float32)
        mats = np.zeros((n_ticks, J, 16), np.float32)
        phases = np.zeros(n_ticks, np.float32)
        self.lib.ref_tick(h, n_ticks, C.c_float(dt), _f(ph), _f(mats), _f(phases))
        return mats, phases

    def blend_weights(self, skeleton_url, clip_url, map_url, mask0_url, mask1_url, w0, w1, J):
        out = np.zeros((2, J), np.float32)
        self.lib.ref_blend_weights.argtypes = [C.c_char_p] * 5 + [C.c_float, C.c_float, _fp]
        self.lib.ref_set_asset_root(self.root.encode())
        n = self.lib.ref_blend_weights(skeleton_url.encode(), clip_url.encode(), map_url.encode(), mask0_url.encode(), mask1_url.encode(),
                                       C.c_float(w0), C.c_float(w1), _f(out))
        if n != J:
            raise RuntimeError("ref_blend_weights failed")
        return out

    def bench_pose(self, h, n_threads, per_thread, full_tick=True):
        return float(self.lib.ref_bench_pose(h, n_threads, per_thread, int(full_tick), None))

    def lbs(self, mesh, palette16, normals=True):
        V = mesh.V
        pos, nrm, tan = (np.ascontiguousarray(a, np.float32) for a in (mesh.pos, mesh.nrm, mesh.tan))
        idx = np.ascontiguousarray(mesh.bind_idx, np.int32)
        w = np.ascontiguousarray(mesh.bind_w, np.float32)
        pal = np.ascontiguousarray(palette16, np.float32)
        op = np.zeros((V, 3), np.float32)
        on = np.zeros((V, 3), np.float32) if normals else None
        ot = np.zeros((V, 3), np.float32) if normals else None
        self.lib.ref_lbs(V, _f(pos), _f(nrm), _f(tan), _i(idx), _f(w), _f(pal), _f(op), _f(on), _f(ot))
        return (op, on, ot) if normals else op

    def bench_lbs(self, mesh, palettes, n_threads):
        """palettes: [n_inst, n_slots, 16].  Returns instances/s."""
        pos, nrm, tan = (np.ascontiguousarray(a, np.float32) for a in (mesh.pos, mesh.nrm, mesh.tan))
        idx = np.ascontiguousarray(mesh.bind_idx, np.int32)
        w = np.ascontiguousarray(mesh.bind_w, np.float32)
        pal = np.ascontiguousarray(palettes, np.float32)
        return float(self.lib.ref_bench_lbs(mesh.V, _f(pos), _f(nrm), _f(tan), _i(idx), _f(w), pal.shape[1], _f(pal),
                                            pal.shape[0], n_threads, None))


def write_player_json(dirpath, skel, clip, amap, mask, sparse=False, prefix=""):
    """Emit the four asset JSON files the reference loader (and ours) reads; returns their urls (relative to dirpath)."""
    from piccolo_b200 import assets as A
    urls = dict(skeleton=prefix + "skeleton.skeleton.json", clip=prefix + "clip.animation_clip.json",
                map=prefix + "anim.skeleton_map.json", mask=prefix + "anim.skeleton_mask.json")
    mask = A.BoneBlendMask(urls["skeleton"], mask.enabled)  # must name the skeleton by the very same url
    A.write_json(A.skeleton_to_json(skel, sparse), os.path.join(dirpath, urls["skeleton"]))
    A.write_json(A.clip_to_json(clip, sparse), os.path.join(dirpath, urls["clip"]))
    A.write_json(A.map_to_json(amap), os.path.join(dirpath, urls["map"]))
    A.write_json(A.mask_to_json(mask), os.path.join(dirpath, urls["mask"]))
    return urls
