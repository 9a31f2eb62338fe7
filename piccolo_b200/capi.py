"""ctypes binding of the C ABI (include/piccolo_b200.h) for the Python test / bench harness.

This is the same boundary a cgo / JNI / C++ caller binds.  There is no fallback: if the
CUDA library is missing, import fails; if no GPU is present, Context() raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PICCOLO_B200_LIB") or os.path.join(PKG, "libpiccolo_b200.so")  # env override: kernel experiments only

PCL_ADVANCE_FIRST_ONLY = 1
PCL_ADVANCE_ALL = 2
PCL_SKIN_POS = 1
PCL_SKIN_POS_NRM_TAN = 7
PCL_STATIC_EXPORT_MAX_BYTES = 8 << 20

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)
_vp = C.c_void_p
_i64 = C.c_int64

# every symbol include/piccolo_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "pcl_version": (C.c_int, []),
    "pcl_device_count": (C.c_int, []),
    "pcl_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "pcl_destroy": (C.c_int, [_vp]),
    "pcl_last_error": (C.c_char_p, [_vp]),
    "pcl_sync": (C.c_int, [_vp]),
    "pcl_launch_count": (_i64, [_vp]),
    "pcl_host_alloc": (C.c_int, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "pcl_host_free": (C.c_int, [_vp, _vp]),
    "pcl_timer_begin": (C.c_int, [_vp]),
    "pcl_timer_end": (C.c_int, [_vp, _fp]),
    "pcl_upload_skeleton": (C.c_int, [_vp, C.c_int32, _ip, _ip, _fp, _fp, _fp, _fp, C.c_int32, C.c_int32, _ip]),
    "pcl_skeleton_bone_count": (C.c_int, [_vp, C.c_int32, _ip]),
    "pcl_skeleton_indices": (C.c_int, [_vp, C.c_int32, _ip]),
    "pcl_skeleton_parents": (C.c_int, [_vp, C.c_int32, _ip]),
    "pcl_upload_clip": (C.c_int, [_vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _ip, _fp, _fp, _fp, C.c_int32, _ip, _ip]),
    "pcl_upload_mask": (C.c_int, [_vp, C.c_int32, C.c_int32, _ip, _ip]),
    "pcl_upload_mesh": (C.c_int, [_vp, C.c_int32, _fp, _fp, _fp, _ip, _fp, C.c_int32, _ip]),
    "pcl_clip_release": (C.c_int, [_vp, C.c_int32]),
    "pcl_mesh_release": (C.c_int, [_vp, C.c_int32]),
    "pcl_skeleton_release": (C.c_int, [_vp, C.c_int32]),
    "pcl_batch_create": (C.c_int, [_vp, C.c_int32, _i64, C.c_int32, C.POINTER(_vp)]),
    "pcl_batch_destroy": (C.c_int, [_vp]),
    "pcl_batch_set": (C.c_int, [_vp, _i64, _i64, _ip, _fp, _fp, _fp, _ip]),
    "pcl_batch_set_frame_async": (C.c_int, [_vp, _i64, _i64, _fp, _fp]),
    "pcl_batch_get_phases": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_batch_advance": (C.c_int, [_vp, C.c_float, C.c_int32]),
    "pcl_batch_pose": (C.c_int, [_vp]),
    "pcl_batch_tick": (C.c_int, [_vp, C.c_float, C.c_int32]),
    "pcl_batch_frame": (C.c_int, [_vp, _ip, _fp, _fp, _ip, _fp]),
    "pcl_batch_frame_result": (C.c_int, [_vp, _fp]),
    "pcl_batch_keep_derived": (C.c_int, [_vp, C.c_int32]),
    "pcl_skinbuf_create": (C.c_int, [_vp, C.c_int32, _i64, C.c_int32, C.POINTER(_vp)]),
    "pcl_skinbuf_destroy": (C.c_int, [_vp]),
    "pcl_batch_skin": (C.c_int, [_vp, C.c_int32, _vp, _i64, _i64, _i64]),
    "pcl_batch_download_palette": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_batch_download_palette_async": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_batch_download_palette3x4_async": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_batch_download_palette3x4": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_batch_download_derived": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_batch_palette_dev": (C.c_int, [_vp, C.POINTER(_vp), _ip, _ip]),
    "pcl_skinbuf_dev": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), _ip]),
    "pcl_skinbuf_download": (C.c_int, [_vp, _i64, _i64, _fp, _fp]),
    "pcl_set_option": (C.c_int, [_vp, C.c_char_p, C.c_int32]),
    "pcl_renderbuf_create": (C.c_int, [_vp, _i64, C.POINTER(_vp)]),
    "pcl_renderbuf_destroy": (C.c_int, [_vp]),
    "pcl_batch_export_render": (C.c_int, [_vp, _i64, _i64, _vp, _i64]),
    "pcl_renderbuf_dev": (C.c_int, [_vp, C.POINTER(_vp)]),
    "pcl_renderbuf_download": (C.c_int, [_vp, _i64, _fp]),
    "pcl_batch_download_palette_static": (C.c_int, [_vp, _i64, _i64, _fp]),
    "pcl_pool_create": (C.c_int, [_vp, C.c_int32, _i64, C.POINTER(_vp)]),
    "pcl_pool_export": (C.c_int, [_vp, C.c_char_p]),
    "pcl_pool_open": (C.c_int, [_vp, C.c_int32, _i64, C.c_char_p, C.POINTER(_vp)]),
    "pcl_pool_destroy": (C.c_int, [_vp]),
    "pcl_batch_attach_pool": (C.c_int, [_vp, _vp, _i64]),
    "pcl_batch_adopt_palettes": (C.c_int, [_vp]),
    "pcl_batch_push_to_pool": (C.c_int, [_vp, _vp, _i64]),
}
PCL_POOL_HANDLE_BYTES = 64

_lib = None


def lib():
    """Load libpiccolo_b200.so (built in-tree by piccolo_b200/build.py).  Missing library => hard error."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s not built: run `python piccolo_b200/build.py` (there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


class PclError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pcl error %d: %s" % (code, msg))
        self.code = code


def _f(a):
    return a.ctypes.data_as(_fp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def _c32(a, dt):
    return None if a is None else np.ascontiguousarray(a, dt)


class Context:
    def __init__(self, device: int = 0):
        self.L = lib()
        h = _vp()
        rc = self.L.pcl_create(device, C.byref(h))
        if rc != 0:
            raise PclError(rc, (self.L.pcl_last_error(None) or b"").decode())
        self.h = h
        self.device = device

    def _ck(self, rc):
        if rc != 0:
            raise PclError(rc, (self.L.pcl_last_error(self.h) or b"").decode())

    def close(self):
        if self.h:
            self.L.pcl_destroy(self.h)
            self.h = None

    def sync(self):
        self._ck(self.L.pcl_sync(self.h))

    def launch_count(self) -> int:
        return int(self.L.pcl_launch_count(self.h))

    def set_option(self, key: str, value: int):
        self._ck(self.L.pcl_set_option(self.h, key.encode(), int(value)))

    def timer_begin(self):
        self._ck(self.L.pcl_timer_begin(self.h))

    def timer_end(self) -> float:
        ms = C.c_float()
        self._ck(self.L.pcl_timer_end(self.h, C.byref(ms)))
        return float(ms.value)

    def pinned(self, shape, dtype=np.float32) -> np.ndarray:
        """numpy view of cudaHostAlloc'd memory (lives as long as the context)."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = _vp()
        self._ck(self.L.pcl_host_alloc(self.h, max(n, 1), C.byref(p)))
        buf = (C.c_char * max(n, 1)).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    # ---- assets ---------------------------------------------------------------------------
    def upload_skeleton(self, s) -> int:
        out = C.c_int32()
        a = [_c32(s.index, np.int32), _c32(s.parent_index, np.int32), _c32(s.bind_pos, np.float32),
             _c32(s.bind_rot, np.float32), _c32(s.bind_scl, np.float32), _c32(s.tpose, np.float32)]
        self._ck(self.L.pcl_upload_skeleton(self.h, s.J, _i(a[0]), _i(a[1]), _f(a[2]), _f(a[3]), _f(a[4]), _f(a[5]),
                                            int(s.is_flat), int(s.in_topological_order), C.byref(out)))
        return out.value

    def skeleton_indices(self, skel: int) -> np.ndarray:
        J = C.c_int32()
        self._ck(self.L.pcl_skeleton_bone_count(self.h, skel, C.byref(J)))
        out = np.zeros(J.value, np.int32)
        self._ck(self.L.pcl_skeleton_indices(self.h, skel, _i(out)))
        return out

    def skeleton_parents(self, skel: int) -> np.ndarray:
        J = C.c_int32()
        self._ck(self.L.pcl_skeleton_bone_count(self.h, skel, C.byref(J)))
        out = np.zeros(J.value, np.int32)
        self._ck(self.L.pcl_skeleton_parents(self.h, skel, _i(out)))
        return out

    def upload_clip(self, skel: int, clip, amap) -> int:
        counts = np.zeros((max(clip.C, 1), 3), np.int32)
        for n in range(clip.C):
            counts[n] = (clip.pos_keys[n].shape[0], clip.rot_keys[n].shape[0], clip.scl_keys[n].shape[0])
        cat = lambda lst, w: (np.ascontiguousarray(np.concatenate([k.reshape(-1, w) for k in lst], 0), np.float32)
                              if lst else np.zeros((1, w), np.float32))
        pos, rot, scl = cat(clip.pos_keys, 3), cat(clip.rot_keys, 4), cat(clip.scl_keys, 3)
        conv = np.ascontiguousarray(amap.convert, np.int32)
        out = C.c_int32()
        self._ck(self.L.pcl_upload_clip(self.h, skel, clip.total_frame, clip.node_count, clip.C, _i(counts), _f(pos), _f(rot),
                                        _f(scl), len(conv), _i(conv), C.byref(out)))
        return out.value

    def upload_mask(self, skel: int, enabled) -> int:
        e = np.ascontiguousarray(enabled, np.int32)
        out = C.c_int32()
        self._ck(self.L.pcl_upload_mask(self.h, skel, len(e), _i(e), C.byref(out)))
        return out.value

    def upload_mesh(self, mesh, max_slot: int) -> int:
        a = [_c32(mesh.pos, np.float32), _c32(mesh.nrm, np.float32), _c32(mesh.tan, np.float32), _c32(mesh.bind_idx, np.int32),
             _c32(mesh.bind_w, np.float32)]
        out = C.c_int32()
        self._ck(self.L.pcl_upload_mesh(self.h, mesh.V, _f(a[0]), _f(a[1]), _f(a[2]), _i(a[3]), _f(a[4]), max_slot, C.byref(out)))
        return out.value

    def release_clip(self, clip: int):
        self._ck(self.L.pcl_clip_release(self.h, clip))

    def release_mesh(self, mesh: int):
        self._ck(self.L.pcl_mesh_release(self.h, mesh))

    def release_skeleton(self, skel: int):
        self._ck(self.L.pcl_skeleton_release(self.h, skel))

    def batch(self, skel: int, N: int, K: int = 1) -> "Batch":
        return Batch(self, skel, N, K)

    def pool(self, skel: int, capacity: int, handle: bytes = None) -> "Pool":
        return Pool(self, skel, capacity, handle)

    def skinbuf(self, mesh: int, V: int, capacity: int, flags: int = PCL_SKIN_POS_NRM_TAN) -> "SkinBuf":
        return SkinBuf(self, mesh, V, capacity, flags)


class Batch:
    def __init__(self, ctx: Context, skel: int, N: int, K: int):
        self.ctx, self.skel, self.N, self.K = ctx, skel, int(N), int(K)
        J = C.c_int32()
        ctx._ck(ctx.L.pcl_skeleton_bone_count(ctx.h, skel, C.byref(J)))
        self.J = J.value
        h = _vp()
        ctx._ck(ctx.L.pcl_batch_create(ctx.h, skel, self.N, self.K, C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.ctx.L.pcl_batch_destroy(self.h)
            self.h = None

    def set(self, first=0, clip_ids=None, phase=None, weight=None, clip_len=None, mask_ids=None):
        arrs = [_c32(clip_ids, np.int32), _c32(phase, np.float32), _c32(weight, np.float32), _c32(clip_len, np.float32),
                _c32(mask_ids, np.int32)]
        count = next(a.size // self.K for a in arrs if a is not None)
        for a in arrs:
            assert a is None or a.size == count * self.K, "arrays must be [count][K]"
        self.ctx._ck(self.ctx.L.pcl_batch_set(self.h, first, count, _i(arrs[0]), _f(arrs[1]), _f(arrs[2]), _f(arrs[3]), _i(arrs[4])))

    def set_frame_async(self, phase=None, weight=None, first=0):
        """Per-frame inputs without a host stall (pcl_batch_set_frame_async)."""
        arrs = [_c32(phase, np.float32), _c32(weight, np.float32)]
        count = next(a.size // self.K for a in arrs if a is not None)
        self.ctx._ck(self.ctx.L.pcl_batch_set_frame_async(self.h, first, count, _f(arrs[0]), _f(arrs[1])))

    def phases(self, first=0, count=None) -> np.ndarray:
        count = self.N - first if count is None else count
        out = np.zeros((count, self.K), np.float32)
        self.ctx._ck(self.ctx.L.pcl_batch_get_phases(self.h, first, count, _f(out)))
        return out

    def advance(self, dt, mode=PCL_ADVANCE_FIRST_ONLY):
        self.ctx._ck(self.ctx.L.pcl_batch_advance(self.h, C.c_float(dt), mode))

    def pose(self):
        self.ctx._ck(self.ctx.L.pcl_batch_pose(self.h))

    def tick(self, dt, mode=PCL_ADVANCE_FIRST_ONLY):
        self.ctx._ck(self.ctx.L.pcl_batch_tick(self.h, C.c_float(dt), mode))

    def frame(self, clip_ids=None, phase=None, weight=None, mask_ids=None, wait=True):
        """pcl_batch_frame: set inputs + pose + read the AnimationResult rows back, one launch and one sync (small batches).
        Returns [N][J][12] (wait=True) or None; frame_result() fetches it later."""
        arrs = [_c32(clip_ids, np.int32), _c32(phase, np.float32), _c32(weight, np.float32), _c32(mask_ids, np.int32)]
        for a in arrs:
            assert a is None or a.size == self.N * self.K, "arrays must be [N][K]"
        out = np.zeros((self.N, self.J, 12), np.float32) if wait else None
        self.ctx._ck(self.ctx.L.pcl_batch_frame(self.h, _i(arrs[0]), _f(arrs[1]), _f(arrs[2]), _i(arrs[3]), _f(out)))
        return out

    def frame_result(self) -> np.ndarray:
        out = np.zeros((self.N, self.J, 12), np.float32)
        self.ctx._ck(self.ctx.L.pcl_batch_frame_result(self.h, _f(out)))
        return out

    def attach_pool(self, pool: "Pool", first: int = 0):
        self.ctx._ck(self.ctx.L.pcl_batch_attach_pool(self.h, pool.h, int(first)))

    def push_to_pool(self, pool: "Pool", first: int = 0):
        self.ctx._ck(self.ctx.L.pcl_batch_push_to_pool(self.h, pool.h, int(first)))

    def adopt_palettes(self):
        self.ctx._ck(self.ctx.L.pcl_batch_adopt_palettes(self.h))

    def keep_derived(self, enable=True):
        self.ctx._ck(self.ctx.L.pcl_batch_keep_derived(self.h, int(enable)))

    def skin(self, mesh: int, out: "SkinBuf", first=0, count=None, slot_first=0):
        count = self.N - first if count is None else count
        self.ctx._ck(self.ctx.L.pcl_batch_skin(self.h, mesh, out.h, first, count, slot_first))

    def download_palette(self, first=0, count=None, out=None, asynchronous=False) -> np.ndarray:
        count = self.N - first if count is None else count
        if out is None:
            out = np.zeros((count, self.J, 16), np.float32)
        fn = self.ctx.L.pcl_batch_download_palette_async if asynchronous else self.ctx.L.pcl_batch_download_palette
        self.ctx._ck(fn(self.h, first, count, _f(out)))
        return out

    def download_palette3x4(self, first=0, count=None, out=None, asynchronous=False) -> np.ndarray:
        """Rows 0..2 of every joint matrix, [count][J][12] (row 3 is the constant 0 0 0 1: expand_3x4 adds it)."""
        count = self.N - first if count is None else count
        if out is None:
            out = np.zeros((count, self.J, 12), np.float32)
        fn = self.ctx.L.pcl_batch_download_palette3x4_async if asynchronous else self.ctx.L.pcl_batch_download_palette3x4
        self.ctx._ck(fn(self.h, first, count, _f(out)))
        return out

    def download_palette_static(self, first=0, count=None) -> np.ndarray:
        """pcl_batch_download_palette_static: same matrices, staged through the library's fixed device array
        (no device allocation; at most PCL_STATIC_EXPORT_MAX_BYTES per call)."""
        count = self.N - first if count is None else count
        out = np.zeros((count, self.J, 16), np.float32)
        self.ctx._ck(self.ctx.L.pcl_batch_download_palette_static(self.h, first, count, _f(out)))
        return out

    def download_derived(self, first=0, count=None) -> np.ndarray:
        count = self.N - first if count is None else count
        out = np.zeros((count, self.J, 10), np.float32)
        self.ctx._ck(self.ctx.L.pcl_batch_download_derived(self.h, first, count, _f(out)))
        return out

    def export_render(self, rb: "RenderBuf", first=0, count=None, draw_first=0):
        count = self.N - first if count is None else count
        self.ctx._ck(self.ctx.L.pcl_batch_export_render(self.h, first, count, rb.h, draw_first))

    def palette_dev(self):
        p, pitch, fps = _vp(), C.c_int32(), C.c_int32()
        self.ctx._ck(self.ctx.L.pcl_batch_palette_dev(self.h, C.byref(p), C.byref(pitch), C.byref(fps)))
        return p.value, pitch.value, fps.value


def expand_3x4(m12: np.ndarray) -> np.ndarray:
    """[..., 12] rows 0..2 -> [..., 16] AnimationResult matrices (row 3 = 0 0 0 1, core/math/matrix4.h:98-119)."""
    out = np.zeros(m12.shape[:-1] + (16,), np.float32)
    out[..., :12] = m12
    out[..., 15] = 1.0
    return out


class Pool:
    """Palette pool (pcl_pool_*): the palettes of several batches - on this or on other GPUs - gathered in one device buffer."""

    def __init__(self, ctx: Context, skel: int, capacity: int, handle: bytes = None):
        self.ctx, self.skel, self.capacity = ctx, skel, int(capacity)
        h = _vp()
        if handle is None:
            ctx._ck(ctx.L.pcl_pool_create(ctx.h, skel, self.capacity, C.byref(h)))
        else:
            assert len(handle) == PCL_POOL_HANDLE_BYTES
            ctx._ck(ctx.L.pcl_pool_open(ctx.h, skel, self.capacity, bytes(handle), C.byref(h)))
        self.h = h

    def export(self) -> bytes:
        buf = C.create_string_buffer(PCL_POOL_HANDLE_BYTES)
        self.ctx._ck(self.ctx.L.pcl_pool_export(self.h, buf))
        return buf.raw

    def close(self):
        if self.h:
            self.ctx._ck(self.ctx.L.pcl_pool_destroy(self.h))
            self.h = None


class SkinBuf:
    def __init__(self, ctx: Context, mesh: int, V: int, capacity: int, flags: int):
        self.ctx, self.mesh, self.V, self.capacity, self.flags = ctx, mesh, int(V), int(capacity), flags
        h = _vp()
        ctx._ck(ctx.L.pcl_skinbuf_create(ctx.h, mesh, self.capacity, flags, C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.ctx.L.pcl_skinbuf_destroy(self.h)
            self.h = None

    def download(self, slot_first=0, count=None):
        count = self.capacity - slot_first if count is None else count
        pos = np.zeros((count, self.V, 3), np.float32)
        nt = np.zeros((count, self.V, 6), np.float32) if self.flags == PCL_SKIN_POS_NRM_TAN else None
        self.ctx._ck(self.ctx.L.pcl_skinbuf_download(self.h, slot_first, count, _f(pos), _f(nt)))
        return pos, nt


class RenderBuf:
    """Per-draw-call joint SSBOs in the reference renderer's layout (pcl_renderbuf_*)."""

    def __init__(self, ctx: Context, draws: int):
        self.ctx, self.draws = ctx, int(draws)
        h = _vp()
        ctx._ck(ctx.L.pcl_renderbuf_create(ctx.h, self.draws, C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.ctx.L.pcl_renderbuf_destroy(self.h)
            self.h = None

    def download(self, draw: int) -> np.ndarray:
        out = np.zeros((64, 1024, 16), np.float32)
        self.ctx._ck(self.ctx.L.pcl_renderbuf_download(self.h, draw, _f(out)))
        return out
