"""ctypes wrapper over oracle/libpm_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (pmtech_b200/) never does.

Pinning status (what pmtech's saved scenes pin, what stays declared): see oracle/pm_oracle.h.

Arrays are numpy, laid out exactly as the reference's component arrays:
  parents u32[N]; entities u64[N]; state_flags u64[N]; local/world f32[N,16];
  bounding_volumes f32[N,13] (min3,max3,tmin3,tmax3,radius — ecs_scene.h:248-255);
  pos_extent f32[N,8] (pos4, extent4 — ecs_scene.h:257-261); frusta f32[V,60]
  (n[6][3], p[6][3], corners[2][4][3] — camera.h:41-46).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpm_oracle.so")

FRUSTUM_FLOATS = 60  # 18 + 18 + 24


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc, -ffp-contract=off)."""
    src = os.path.join(_HERE, "pm_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(
        os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, "pm_oracle.h"))
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


class _Scene(C.Structure):
    _fields_ = [
        ("num_entities", C.c_size_t),
        ("entities", C.c_void_p),
        ("state_flags", C.c_void_p),
        ("parents", C.c_void_p),
        ("transforms", C.c_void_p),
        ("local_matrices", C.c_void_p),
        ("world_matrices", C.c_void_p),
        ("bounding_volumes", C.c_void_p),
        ("pos_extent", C.c_void_p),
        ("renderable_extents", C.c_float * 6),
        ("has_rb_matrix", C.c_void_p),
    ]


class _Camera(C.Structure):
    _fields_ = [
        ("pos", C.c_float * 3),
        ("view", C.c_float * 16),
        ("proj", C.c_float * 16),
        ("camera_frustum", C.c_float * FRUSTUM_FLOATS),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_update_scene.argtypes = [C.POINTER(_Scene)]
        L.orc_update_transforms.argtypes = [C.POINTER(_Scene)]
        L.orc_update_bounds.argtypes = [C.POINTER(_Scene)]
        L.orc_expand_parent_extents.argtypes = [C.POINTER(_Scene)]
        L.orc_update_draw_calls.argtypes = [C.POINTER(_Scene), C.c_void_p, C.c_void_p, C.c_float]
        L.orc_inverse4x4.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_apply_root_motion.argtypes = [C.POINTER(_Scene), C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_skin_palettes.argtypes = [C.POINTER(_Scene), C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_sb_count.argtypes = [C.c_void_p]
        L.orc_sb_count.restype = C.c_int
        L.orc_sb_free.argtypes = [C.c_void_p]
        L.orc_sb_push.argtypes = [C.POINTER(C.c_void_p), C.c_uint32]
        L.orc_sb_from_array.argtypes = [C.c_void_p, C.c_uint32]
        L.orc_sb_from_array.restype = C.c_void_p
        L.orc_filter_entities_scalar.argtypes = [C.POINTER(_Scene), C.POINTER(C.c_void_p)]
        for f in (L.orc_frustum_cull_aabb_scalar, L.orc_frustum_cull_sphere_scalar):
            f.argtypes = [C.POINTER(_Scene), C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        L.orc_frame.argtypes = [C.POINTER(_Scene), C.c_void_p, C.c_uint32, C.c_int, C.POINTER(C.c_void_p)]
        L.orc_frame_omp.argtypes = L.orc_frame.argtypes
        L.orc_frame_omp.restype = C.c_int
        L.orc_mat4_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_transform_vector.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_quat_get_matrix.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_quat_expand.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_plane_distance.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_plane_distance.restype = C.c_float
        L.orc_camera_create_perspective.argtypes = [C.POINTER(_Camera)] + [C.c_float] * 4
        L.orc_camera_create_orthographic.argtypes = [C.POINTER(_Camera)] + [C.c_float] * 6
        L.orc_camera_create_cubemap.argtypes = [C.POINTER(_Camera), C.c_float, C.c_float]
        L.orc_camera_set_cubemap_face.argtypes = [C.POINTER(_Camera), C.c_uint32]
        L.orc_camera_look_at.argtypes = [C.POINTER(_Camera), C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_camera_update_frustum.argtypes = [C.POINTER(_Camera)]
        L.orc_camera_spot_shadow.argtypes = [C.POINTER(_Camera), C.c_void_p]
        L.orc_camera_update_shadow_frustum.argtypes = [C.POINTER(_Camera), C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def set_num_threads(n=None):
    """threads of the OpenMP variant; default = the cores this process may run on (torchrun forces OMP_NUM_THREADS=1)"""
    n = len(os.sched_getaffinity(0)) if n is None else int(n)
    lib().orc_set_num_threads(n)
    return n


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _req(a, dtype, shape_tail=None):
    a = np.ascontiguousarray(a, dtype=dtype)
    if shape_tail is not None:
        assert a.shape[1:] == shape_tail, (a.shape, shape_tail)
    return a


class OracleScene:
    """Holds host copies of the component arrays (copies of what it is given) and runs the
    reference loops over them."""

    def __init__(self, parents, local_matrices, entities, state_flags, bounding_volumes,
                 world_matrices=None, pos_extent=None, has_rb_matrix=None, transforms=None):
        n = len(parents)
        self.n = n
        self.parents = _req(parents, np.uint32).copy()
        self.local = _req(local_matrices, np.float32, (16,)).copy()
        self.entities = _req(entities, np.uint64).copy()
        self.state_flags = _req(state_flags, np.uint64).copy()
        self.bv = _req(bounding_volumes, np.float32, (13,)).copy()
        self.world = (np.zeros((n, 16), np.float32) if world_matrices is None
                      else _req(world_matrices, np.float32, (16,)).copy())
        self.pos_extent = (np.zeros((n, 8), np.float32) if pos_extent is None
                           else _req(pos_extent, np.float32, (8,)).copy())
        self.has_rb = None if has_rb_matrix is None else _req(has_rb_matrix, np.uint8).copy()
        self.transforms = None if transforms is None else _req(transforms, np.float32, (10,)).copy()
        self._s = _Scene()
        self._bind()

    def _bind(self):
        s = self._s
        s.num_entities = self.n
        s.entities = _ptr(self.entities)
        s.state_flags = _ptr(self.state_flags)
        s.parents = _ptr(self.parents)
        s.transforms = _ptr(self.transforms)
        s.local_matrices = _ptr(self.local)
        s.world_matrices = _ptr(self.world)
        s.bounding_volumes = _ptr(self.bv)
        s.pos_extent = _ptr(self.pos_extent)
        s.has_rb_matrix = _ptr(self.has_rb)

    @property
    def renderable_extents(self):
        return np.array(list(self._s.renderable_extents), dtype=np.float32)

    # --- update_scene (ecs_scene.cpp:1405-1556) ---
    def update_scene(self):
        lib().orc_update_scene(C.byref(self._s))

    def update_transforms(self):
        lib().orc_update_transforms(C.byref(self._s))

    def update_bounds(self):
        lib().orc_update_bounds(C.byref(self._s))

    def expand_parent_extents(self):
        lib().orc_expand_parent_extents(C.byref(self._s))

    # --- animation hand-off (ecs_scene.cpp:1286-1304, :1917-1959) ---
    def apply_root_motion(self, targets, rotations, deltas):
        t = _req(targets, np.uint32)
        q = _req(np.asarray(rotations, np.float32).reshape(-1, 4), np.float32, (4,))
        d = _req(np.asarray(deltas, np.float32).reshape(-1, 3), np.float32, (3,))
        lib().orc_apply_root_motion(C.byref(self._s), _ptr(t), _ptr(q), _ptr(d), C.c_uint32(t.shape[0]))

    def skin_palettes(self, skins, bind_matrices, num_palette_matrices):
        sk = np.ascontiguousarray(skins)
        assert sk.dtype.itemsize == 16
        b = _req(np.asarray(bind_matrices, np.float32).reshape(-1, 16), np.float32, (16,))
        out = np.zeros((num_palette_matrices, 16), np.float32)
        lib().orc_skin_palettes(C.byref(self._s), _ptr(sk), C.c_uint32(sk.shape[0]), _ptr(b), _ptr(out))
        return out

    def update_draw_calls(self, cbuffer, draw_call_data, time_ms):
        """ecs_scene.cpp:1962-1991 over draw_call_data ([n, 40] f32 records, updated in place)."""
        cb = _req(cbuffer, np.uint32)
        assert draw_call_data.dtype == np.float32 and draw_call_data.shape == (self.n, 40) and draw_call_data.flags.c_contiguous
        lib().orc_update_draw_calls(C.byref(self._s), _ptr(cb), _ptr(draw_call_data), C.c_float(time_ms))
        return draw_call_data

    # --- cull (ecs_cull.cpp:21-105) ---
    @staticmethod
    def _take(sb):
        L = lib()
        cnt = L.orc_sb_count(sb)
        if cnt == 0:
            out = np.zeros(0, np.uint32)
        else:
            out = np.ctypeslib.as_array(C.cast(sb, C.POINTER(C.c_uint32)), shape=(cnt,)).copy()
        L.orc_sb_free(sb)
        return out

    @staticmethod
    def _make_sb(indices):
        idx = np.ascontiguousarray(indices, dtype=np.uint32)
        return C.c_void_p(lib().orc_sb_from_array(_ptr(idx), idx.size))

    def filter_entities(self):
        sb = C.c_void_p(None)
        lib().orc_filter_entities_scalar(C.byref(self._s), C.byref(sb))
        return self._take(sb)

    def frustum_cull(self, frustum, entities_in, mode="aabb"):
        fr = _req(frustum, np.float32).reshape(-1)
        assert fr.size == FRUSTUM_FLOATS
        sb_in = self._make_sb(entities_in)
        sb_out = C.c_void_p(None)
        f = lib().orc_frustum_cull_aabb_scalar if mode == "aabb" else lib().orc_frustum_cull_sphere_scalar
        f(C.byref(self._s), _ptr(fr), sb_in, C.byref(sb_out))
        lib().orc_sb_free(sb_in)
        return self._take(sb_out)

    def frame(self, frusta, mode="aabb", omp=False):
        """update_scene + per-view filter+cull.  Returns (lists, threads_used)."""
        fr = _req(frusta, np.float32).reshape(-1, FRUSTUM_FLOATS)
        v = fr.shape[0]
        outs = (C.c_void_p * max(v, 1))()
        m = 0 if mode == "aabb" else 1
        if omp:
            threads = lib().orc_frame_omp(C.byref(self._s), _ptr(fr), v, m, outs)
        else:
            lib().orc_frame(C.byref(self._s), _ptr(fr), v, m, outs)
            threads = 1
        return [self._take(C.c_void_p(outs[i])) for i in range(v)], threads


# ---- maths pins -------------------------------------------------------------------------
def mat4_mul(a, b):
    a = _req(a, np.float32).reshape(16)
    b = _req(b, np.float32).reshape(16)
    o = np.zeros(16, np.float32)
    lib().orc_mat4_mul(_ptr(o), _ptr(a), _ptr(b))
    return o


def transform_vector(m, v):
    m = _req(m, np.float32).reshape(16)
    v = _req(v, np.float32).reshape(3)
    o = np.zeros(3, np.float32)
    lib().orc_transform_vector(_ptr(o), _ptr(m), _ptr(v))
    return o


def quat_get_matrix(q):
    """Returns (matrix, q after get_matrix normalised it in place)."""
    q = np.array(q, np.float32).reshape(4)
    o = np.zeros(16, np.float32)
    lib().orc_quat_get_matrix(_ptr(o), _ptr(q))
    return o, q


def quat_expand(q):
    q = np.array(q, np.float32).reshape(4)
    o = np.zeros(16, np.float32)
    lib().orc_quat_expand(_ptr(o), _ptr(q))
    return o


# ---- cameras (camera.cpp) ----------------------------------------------------------------
class OracleCamera:
    def __init__(self):
        self._c = _Camera()

    def create_perspective(self, fov_degrees, aspect, near, far):
        lib().orc_camera_create_perspective(C.byref(self._c), fov_degrees, aspect, near, far)
        return self

    def create_orthographic(self, l, r, b, t, n, f):
        lib().orc_camera_create_orthographic(C.byref(self._c), l, r, b, t, n, f)
        return self

    def create_cubemap(self, near, far):
        lib().orc_camera_create_cubemap(C.byref(self._c), near, far)
        return self

    def set_pos(self, pos):
        for i in range(3):
            self._c.pos[i] = float(pos[i])
        return self

    def set_cubemap_face(self, face):
        lib().orc_camera_set_cubemap_face(C.byref(self._c), face)
        return self

    def look_at(self, pos, at, up=(0.0, 1.0, 0.0)):
        p = np.asarray(pos, np.float32)
        a = np.asarray(at, np.float32)
        u = np.asarray(up, np.float32)
        lib().orc_camera_look_at(C.byref(self._c), _ptr(p), _ptr(a), _ptr(u))
        return self

    def update_frustum(self):
        lib().orc_camera_update_frustum(C.byref(self._c))
        return self

    def spot_shadow(self, world):
        w = np.ascontiguousarray(world, np.float32).reshape(16)
        lib().orc_camera_spot_shadow(C.byref(self._c), _ptr(w))
        return self

    def update_shadow_frustum(self, light_dir, mn, mx):
        d = np.asarray(light_dir, np.float32)
        a = np.asarray(mn, np.float32)
        b = np.asarray(mx, np.float32)
        lib().orc_camera_update_shadow_frustum(C.byref(self._c), _ptr(d), _ptr(a), _ptr(b))
        return self

    @property
    def frustum(self):
        return np.array(list(self._c.camera_frustum), dtype=np.float32)

    @property
    def view(self):
        return np.array(list(self._c.view), dtype=np.float32)

    @property
    def proj(self):
        return np.array(list(self._c.proj), dtype=np.float32)
