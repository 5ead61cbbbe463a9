"""Host-side mirror of the reference interface for this path, over the C ABI.

`EcsScene` carries numpy arrays named and laid out like put::ecs::ecs_scene's component arrays
(core/put/source/ecs/ecs_scene.h:460-543); the module-level functions carry the reference's
names and argument meaning:

    update_scene(scene, dt)                         ecs_scene.h:559  / ecs_scene.cpp:1373
    filter_entities_scalar(scene)                   ecs_cull.h:16    / ecs_cull.cpp:84
    frustum_cull_aabb_scalar(scene, cam, ents_in)   ecs_cull.h:17    / ecs_cull.cpp:21
    frustum_cull_sphere_scalar(scene, cam, ents_in) ecs_cull.h:18    / ecs_cull.cpp:53
    frustum_cull_aabb / frustum_cull_sphere         ecs_cull.h:21-22 (dispatch to scalar, :533-543)

All compute happens in libpmtech_b200.so on the GPU; nothing here evaluates the path on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import (BOUNDING_VOLUMES, CULL_AABB, CULL_FILTER_ONLY, CULL_SPHERE, ENTITIES, FRUSTUM_FLOATS,
                   HAS_RB_MATRIX, LOCAL_MATRICES, PARENTS, POS_EXTENT, STATE_FLAGS, WORLD_MATRICES, PmbError,
                   _check, _ptr)

_MODES = {"aabb": CULL_AABB, "sphere": CULL_SPHERE, "filter": CULL_FILTER_ONLY}


class DeviceScene:
    """RAII wrapper of a pmb_scene handle (device mirrors + kernels)."""

    def __init__(self, capacity: int, device: int = 0):
        self._L = capi.load()
        h = C.c_void_p()
        _check(self._L.pmb_scene_create(C.byref(h), int(capacity), int(device)))
        self._h = h
        self.capacity = int(capacity)
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._L.pmb_scene_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- plumbing ----
    def set_stream(self, cuda_stream_ptr):
        _check(self._L.pmb_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def set_option(self, option, value):
        """pmb_set_option: capi.OPT_HIERARCHY (HIERARCHY_AUTO / _CHAINS / _WAVEFRONT), capi.OPT_FUSED_TAIL (0 / 1)"""
        _check(self._L.pmb_set_option(self._h, int(option), int(value)))

    def download_trace(self, num_tiles):
        out = np.zeros((int(num_tiles), 4), np.uint64)
        _check(self._L.pmb_download_trace(self._h, _ptr(out), int(num_tiles)))
        return out

    def synchronize(self):
        _check(self._L.pmb_synchronize(self._h))

    def resize(self, capacity):
        _check(self._L.pmb_scene_resize(self._h, int(capacity)))
        self.capacity = int(capacity)

    def upload(self, array_id, host, first=0, count=None):
        dt, tail = capi._ARRAY_SPEC[array_id]
        a = np.ascontiguousarray(host, dtype=dt)
        if count is None:
            count = a.shape[0] if a.ndim else 0
        assert a.shape[1:] == tail, (a.shape, tail)
        _check(self._L.pmb_upload(self._h, array_id, _ptr(a), int(first), int(count)))
        return a  # keep alive until the caller synchronises

    def download(self, array_id, first=0, count=None, out=None):
        dt, tail = capi._ARRAY_SPEC[array_id]
        if count is None:
            count = self.capacity - first
        if out is None:
            out = np.empty((count,) + tail, dtype=dt)
        _check(self._L.pmb_download(self._h, array_id, _ptr(out), int(first), int(count)))
        return out

    def device_ptr(self, array_id) -> int:
        return self._L.pmb_device_ptr(self._h, array_id) or 0

    # ---- stages ----
    def update_scene(self, n):
        _check(self._L.pmb_update_scene(self._h, int(n)))

    def renderable_extents(self):
        out = np.empty(6, np.float32)
        _check(self._L.pmb_get_renderable_extents(self._h, _ptr(out)))
        return out

    @staticmethod
    def _views(frusta):
        fr = np.ascontiguousarray(frusta, dtype=np.float32).reshape(-1, FRUSTUM_FLOATS)
        return fr, fr.shape[0]

    def cull_views(self, n, frusta, mode="aabb"):
        fr, nv = self._views(frusta)
        _check(self._L.pmb_cull_views(self._h, int(n), _ptr(fr), nv, _MODES[mode]))
        return nv

    def mark_transforms_dirty(self, first, count):
        """entities[first:first+count] |= e_cmp::transform on the device"""
        _check(self._L.pmb_mark_transforms_dirty(self._h, int(first), int(count)))

    # ---- shadow views fitted on the device (ecs_scene.cpp:538-554, camera.cpp:384-416) ----
    def reserve_views(self, num_views):
        _check(self._L.pmb_reserve_views(self._h, int(num_views)))

    def cull_shadow_views(self, n, light_directions, constraints=None, mode="aabb", first_view=0):
        d = np.ascontiguousarray(light_directions, np.float32).reshape(-1, 3)
        c = None if constraints is None else np.ascontiguousarray(constraints, np.float32).reshape(6)
        _check(self._L.pmb_cull_shadow_views(self._h, int(n), _ptr(d), d.shape[0], _ptr(c), _MODES[mode], int(first_view)))
        return d.shape[0]

    def cull_spot_shadow_views(self, n, light_entities, mode="aabb", first_view=0):
        e = np.ascontiguousarray(light_entities, np.uint32).reshape(-1)
        _check(self._L.pmb_cull_spot_shadow_views(self._h, int(n), _ptr(e), e.shape[0], _MODES[mode], int(first_view)))
        return e.shape[0]

    def download_shadow_cameras(self, count):
        """[count] records of pmb_camera as float32 rows: pos 0:3, view 3:19, proj 19:35, frustum 35:95"""
        out = np.zeros((count, 3 + 16 + 16 + 60), np.float32)
        if count:
            _check(self._L.pmb_download_shadow_cameras(self._h, _ptr(out), int(count)))
        return out

    def frame(self, n, frusta, mode="aabb"):
        fr, nv = self._views(frusta)
        _check(self._L.pmb_frame(self._h, int(n), _ptr(fr), nv, _MODES[mode]))
        return nv

    def counts(self, nv):
        out = np.zeros(nv, np.uint32)
        if nv:
            _check(self._L.pmb_get_counts(self._h, _ptr(out), nv))
        return out

    def lists(self, nv):
        cnt = self.counts(nv)
        res = []
        for v in range(nv):
            buf = np.empty(int(cnt[v]), np.uint32)
            if cnt[v]:
                _check(self._L.pmb_get_list(self._h, v, _ptr(buf), int(cnt[v])))
            res.append(buf)
        return res

    def lists_sb(self, nv):
        """the stretchy-buffer form: returns (numpy copies, raw header ints) and frees the buffers"""
        ptrs = (C.c_void_p * max(nv, 1))()
        _check(self._L.pmb_get_lists_sb(self._h, ptrs, nv))
        libc = C.CDLL(None)
        libc.free.argtypes = [C.c_void_p]
        res = []
        for v in range(nv):
            if not ptrs[v]:
                res.append((np.zeros(0, np.uint32), None))
                continue
            hdr = np.ctypeslib.as_array(C.cast(ptrs[v] - 8, C.POINTER(C.c_int)), shape=(2,)).copy()
            data = np.ctypeslib.as_array(C.cast(ptrs[v], C.POINTER(C.c_uint32)), shape=(int(hdr[1]),)).copy()
            libc.free(C.c_void_p(ptrs[v] - 8))  # sb_free: free(stb__sbraw(a)), data_struct.h:22,30
            res.append((data, hdr))
        return res

    def frame_host(self, n, frusta, mode="aabb", *, parents=None, local_matrices=None, entities=None,
                   state_flags=None, bounding_volumes=None, has_rb_matrix=None, transforms=None, world_matrices_out=None,
                   bounding_volumes_out=None, pos_extent_out=None, renderable_extents_out=None, lists_out=None,
                   counts_out=None):
        """pmb_frame_host: H2D of the given arrays + frame + D2H of lists/outputs, one call."""
        fr, nv = self._views(frusta)
        io = capi.HostArrays()
        for name, arr in (("parents", parents), ("local_matrices", local_matrices), ("entities", entities),
                          ("state_flags", state_flags), ("bounding_volumes", bounding_volumes),
                          ("has_rb_matrix", has_rb_matrix), ("transforms", transforms), ("world_matrices_out", world_matrices_out),
                          ("bounding_volumes_out", bounding_volumes_out), ("pos_extent_out", pos_extent_out),
                          ("renderable_extents_out", renderable_extents_out)):
            setattr(io, name, None if arr is None else arr.ctypes.data)
        lp = (C.c_void_p * max(nv, 1))()
        if lists_out is not None:
            for v in range(nv):
                lp[v] = lists_out[v].ctypes.data
        if counts_out is None:
            counts_out = np.zeros(max(nv, 1), np.uint32)
        _check(self._L.pmb_frame_host(self._h, int(n), C.byref(io), _ptr(fr), nv, _MODES[mode], lp, _ptr(counts_out)))
        return counts_out[:nv]

    def update_draw_calls(self, n, time_ms=0.0):
        """ecs_scene.cpp:1962-1991 into the PMB_DRAW_CALLS mirror"""
        _check(self._L.pmb_update_draw_calls(self._h, int(n), float(time_ms)))

    # ---- animation hand-off (ecs_scene.cpp:1286-1304, :1917-1959) ----
    def apply_root_motion(self, targets, rotations, deltas, sync=True):
        """sync=False: the caller keeps the three arrays alive until the stream has consumed them"""
        t = np.ascontiguousarray(targets, np.uint32)
        q = np.ascontiguousarray(rotations, np.float32).reshape(-1, 4)
        d = np.ascontiguousarray(deltas, np.float32).reshape(-1, 3)
        assert t.shape[0] == q.shape[0] == d.shape[0]
        _check(self._L.pmb_apply_root_motion(self._h, _ptr(t), _ptr(q), _ptr(d), t.shape[0]))
        if sync:
            self.synchronize()  # the host arrays may go away

    def set_skins(self, skins, bind_matrices, num_palette_matrices):
        sk = np.ascontiguousarray(skins, capi.SKIN_DTYPE)
        b = np.ascontiguousarray(bind_matrices, np.float32).reshape(-1, 16)
        _check(self._L.pmb_skins_set(self._h, _ptr(sk), sk.shape[0], _ptr(b), b.shape[0], int(num_palette_matrices)))
        self._num_palette = int(num_palette_matrices)

    def skin_palettes(self):
        _check(self._L.pmb_skin_palettes(self._h))

    def download_palettes(self, first=0, count=None):
        if count is None:
            count = self._num_palette - first
        out = np.empty((count, 16), np.float32)
        _check(self._L.pmb_download_palettes(self._h, _ptr(out), int(first), int(count)))
        return out

    # ---- multi-GPU list exchange helpers ----
    def device_list_ptr(self, view) -> int:
        return self._L.pmb_device_list(self._h, int(view)) or 0

    def device_counts_ptr(self) -> int:
        return self._L.pmb_device_counts(self._h) or 0

    def pack_lists(self, counts, add, dst_device_ptr):
        c = np.ascontiguousarray(counts, np.uint32)
        _check(self._L.pmb_pack_lists(self._h, _ptr(c), c.shape[0], int(add), C.c_void_p(dst_device_ptr)))

    def copy_segments(self, src_device_ptr, dst_device_ptr, segments):
        seg = np.ascontiguousarray(segments, capi.SEGMENT_DTYPE)
        _check(self._L.pmb_copy_segments(self._h, C.c_void_p(src_device_ptr), C.c_void_p(dst_device_ptr), _ptr(seg),
                                         seg.shape[0]))

    def pack_row(self, num_views, add, row_device_ptr, row_stride):
        _check(self._L.pmb_pack_row(self._h, int(num_views), int(add), C.c_void_p(row_device_ptr), int(row_stride)))

    def unpack_rows(self, rows_device_ptr, world, row_stride, num_views, dst_device_ptr, dst_capacity, meta_device_ptr,
                    cuda_stream_ptr=None):
        _check(self._L.pmb_unpack_rows(self._h, C.c_void_p(rows_device_ptr), int(world), int(row_stride), int(num_views),
                                       C.c_void_p(dst_device_ptr), int(dst_capacity), C.c_void_p(meta_device_ptr),
                                       C.c_void_p(cuda_stream_ptr or 0)))

    def publish_counts(self, row_device_ptr, num_views, rank, world, peer_table_ptrs, cuda_stream_ptr=None):
        pt = np.ascontiguousarray(peer_table_ptrs, np.uint64)
        _check(self._L.pmb_publish_counts(self._h, C.c_void_p(row_device_ptr), int(num_views), int(rank), int(world), _ptr(pt),
                                          C.c_void_p(cuda_stream_ptr or 0)))

    def push_row(self, row_device_ptr, row_stride, num_views, rank, world, table_device_ptr, peer_list_ptrs, lists_capacity,
                 meta_device_ptr, cuda_stream_ptr=None):
        pl = np.ascontiguousarray(peer_list_ptrs, np.uint64)
        _check(self._L.pmb_push_row(self._h, C.c_void_p(row_device_ptr), int(row_stride), int(num_views), int(rank), int(world),
                                    C.c_void_p(table_device_ptr), _ptr(pl), int(lists_capacity), C.c_void_p(meta_device_ptr),
                                    C.c_void_p(cuda_stream_ptr or 0)))

    # ---- instrumentation ----
    def set_profiling(self, on):
        _check(self._L.pmb_set_profiling(self._h, int(on)))  # 0 off, 1 every launch group, 2 the streaming pass only

    def kernel_stats(self):
        out = {}
        for i, name in enumerate(capi.KERNEL_NAMES):
            ms, n = C.c_double(), C.c_uint64()
            _check(self._L.pmb_kernel_stats(self._h, i, C.byref(ms), C.byref(n)))
            out[name] = (ms.value, n.value)
        return out

    def reset_kernel_stats(self):
        _check(self._L.pmb_reset_kernel_stats(self._h))

    def total_launches(self):
        return int(self._L.pmb_total_launches(self._h))
