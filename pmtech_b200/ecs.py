"""put::ecs-shaped host API (names, argument meaning and list semantics of the reference) on top
of DeviceScene.  See scene.py for the citations.  `EcsScene` owns host numpy arrays laid out as
the reference's cmp_arrays and a device handle; `update_scene` uploads the input arrays, runs
the device stage and reads the outputs back into the host arrays, like a shimmed
put::ecs::update_scene would."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .scene import DeviceScene
from .capi import (BOUNDING_VOLUMES, ENTITIES, HAS_RB_MATRIX, LOCAL_MATRICES, PARENTS, POS_EXTENT, STATE_FLAGS,
                   WORLD_MATRICES)

# e_cmp (ecs_scene.h:144-171) / e_state (ecs_scene.h:69-83) bits used by the path
CMP_ALLOCATED, CMP_GEOMETRY, CMP_PHYSICS, CMP_MATERIAL = 1 << 0, 1 << 1, 1 << 2, 1 << 4
CMP_BONE, CMP_ANIM_CONTROLLER, CMP_TRANSFORM, CMP_SUB_INSTANCE = 1 << 7, 1 << 9, 1 << 12, 1 << 14
STATE_HIDDEN, STATE_SYNC_PHYSICS_TRANSFORM = 1 << 2, 1 << 7


class Camera:
    """put::camera (camera.h:55-73): pos, view, proj, camera_frustum + the producers of camera.cpp."""

    def __init__(self):
        self._c = capi.CameraStruct()
        self._L = capi.load()

    def create_perspective(self, fov_degrees, aspect, near, far):
        self._L.pmb_camera_create_perspective(C.byref(self._c), fov_degrees, aspect, near, far)
        return self

    def create_orthographic(self, l, r, b, t, n, f):
        self._L.pmb_camera_create_orthographic(C.byref(self._c), l, r, b, t, n, f)
        return self

    def create_cubemap(self, near, far):
        self._L.pmb_camera_create_cubemap(C.byref(self._c), near, far)
        return self

    def set_pos(self, pos):
        for i in range(3):
            self._c.pos[i] = float(pos[i])
        return self

    def set_cubemap_face(self, face):
        self._L.pmb_camera_set_cubemap_face(C.byref(self._c), int(face))
        return self

    def look_at(self, pos, at, up=(0.0, 1.0, 0.0)):
        p, a, u = (np.asarray(x, np.float32) for x in (pos, at, up))
        self._L.pmb_camera_look_at(C.byref(self._c), capi._ptr(p), capi._ptr(a), capi._ptr(u))
        return self

    def update_frustum(self):
        self._L.pmb_camera_update_frustum(C.byref(self._c))
        return self

    def spot_shadow(self, world_matrix):
        """shadow_camera_from_entity, spot branch (ecs_scene.cpp:556-570)"""
        w = np.ascontiguousarray(world_matrix, np.float32).reshape(16)
        self._L.pmb_camera_spot_shadow(C.byref(self._c), capi._ptr(w))
        return self

    def update_shadow_frustum(self, light_dir, mn, mx):
        d, a, b = (np.asarray(x, np.float32) for x in (light_dir, mn, mx))
        self._L.pmb_camera_update_shadow_frustum(C.byref(self._c), capi._ptr(d), capi._ptr(a), capi._ptr(b))
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


class EcsScene:
    """The slice of put::ecs::ecs_scene this path touches, plus its device mirror."""

    def __init__(self, soa_size: int, device: int = 0):
        n = int(soa_size)
        self.soa_size = n
        self.num_entities = 0
        # resize_scene_buffers zero-fills (ecs_scene.cpp:122-152)
        self.entities = np.zeros(n, np.uint64)
        self.state_flags = np.zeros(n, np.uint64)
        self.parents = np.zeros(n, np.uint32)
        self.local_matrices = np.zeros((n, 16), np.float32)
        self.world_matrices = np.zeros((n, 16), np.float32)
        self.bounding_volumes = np.zeros((n, 13), np.float32)
        self.pos_extent = np.zeros((n, 8), np.float32)
        self.renderable_extents = np.zeros(6, np.float32)
        self.has_rb_matrix = None
        self.dev = DeviceScene(n, device)
        self._views_cached = None

    def close(self):
        self.dev.close()

    @classmethod
    def from_arrays(cls, arrays: dict, device: int = 0):
        n = len(arrays["parents"])
        s = cls(n, device)
        s.num_entities = int(arrays.get("num_entities", n))
        for k in ("entities", "state_flags", "parents", "local_matrices", "bounding_volumes"):
            getattr(s, k)[...] = arrays[k]
        for k in ("world_matrices", "pos_extent"):
            if arrays.get(k) is not None:
                getattr(s, k)[...] = arrays[k]
        if arrays.get("has_rb_matrix") is not None:
            s.has_rb_matrix = np.ascontiguousarray(arrays["has_rb_matrix"], np.uint8)
        return s

    def upload_inputs(self, include_outputs=False):
        d, n = self.dev, self.num_entities
        d.upload(PARENTS, self.parents[:n])
        d.upload(LOCAL_MATRICES, self.local_matrices[:n])
        d.upload(ENTITIES, self.entities[:n])
        d.upload(STATE_FLAGS, self.state_flags[:n])
        d.upload(BOUNDING_VOLUMES, self.bounding_volumes[:n])
        if self.has_rb_matrix is not None:
            d.upload(HAS_RB_MATRIX, self.has_rb_matrix[:n])
        if include_outputs:
            d.upload(WORLD_MATRICES, self.world_matrices[:n])
            d.upload(POS_EXTENT, self.pos_extent[:n])
        d.synchronize()

    def download_outputs(self):
        d, n = self.dev, self.num_entities
        d.download(WORLD_MATRICES, 0, n, out=self.world_matrices[:n])
        d.download(BOUNDING_VOLUMES, 0, n, out=self.bounding_volumes[:n])
        d.download(POS_EXTENT, 0, n, out=self.pos_extent[:n])
        d.download(ENTITIES, 0, n, out=self.entities[:n])
        d.download(STATE_FLAGS, 0, n, out=self.state_flags[:n])
        self.renderable_extents[...] = d.renderable_extents()


def update_scene(scene: EcsScene, dt: float = 0.0, upload=True, download=True):
    """put::ecs::update_scene loops A-C (ecs_scene.cpp:1405-1556).  `dt` is unused by these loops."""
    if upload:
        scene.upload_inputs(include_outputs=True)
    scene.dev.update_scene(scene.num_entities)
    if download:
        scene.download_outputs()


def filter_entities_scalar(scene: EcsScene) -> np.ndarray:
    """ecs_cull.cpp:84-105 — returns the filtered index list (ascending)."""
    scene.dev.cull_views(scene.num_entities, np.zeros((1, capi.FRUSTUM_FLOATS), np.float32), "filter")
    return scene.dev.lists(1)[0]


def _cull(scene, cam, entities_in, mode):
    fr = cam.frustum if hasattr(cam, "frustum") else np.asarray(cam, np.float32)
    scene.dev.cull_views(scene.num_entities, fr.reshape(1, -1), mode)
    vis = scene.dev.lists(1)[0]
    if entities_in is None:
        return vis
    # the reference culls the list it is handed (normally filter_entities_scalar's output) and keeps
    # its order; the device list is the cull of *all* filtered entities in ascending order
    ents = np.asarray(entities_in, np.uint32)
    keep = np.isin(ents, vis, assume_unique=False)
    return ents[keep]


def frustum_cull_aabb_scalar(scene, cam, entities_in=None):
    """ecs_cull.cpp:21-51"""
    return _cull(scene, cam, entities_in, "aabb")


def frustum_cull_sphere_scalar(scene, cam, entities_in=None):
    """ecs_cull.cpp:53-82"""
    return _cull(scene, cam, entities_in, "sphere")


# ecs_cull.cpp:533-543: the dispatchers are hard-wired to the scalar versions
frustum_cull_aabb = frustum_cull_aabb_scalar
frustum_cull_sphere = frustum_cull_sphere_scalar


def simd_init():
    """ecs_cull.cpp:530-531 — nothing to detect: the device path is the only implementation."""
    capi.load()
