"""ctypes binding of include/pmtech_b200.h (libpmtech_b200.so).

Thin and literal: one Python function per C entry point, numpy arrays in the reference's own
layouts.  There is no fallback: if the shared object is missing `load()` raises, and on a
machine without a CUDA device `Scene(...)` raises PmbError(PMB_ERR_NO_DEVICE).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "lib", "libpmtech_b200.so")

MAX_VIEWS = 16
FRUSTUM_FLOATS = 60  # n[6][3], p[6][3], corners[2][4][3]  (camera.h:41-46)

# pmb_array
PARENTS, LOCAL_MATRICES, WORLD_MATRICES, ENTITIES, STATE_FLAGS, BOUNDING_VOLUMES, POS_EXTENT, HAS_RB_MATRIX, TRANSFORMS, CBUFFER, DRAW_CALLS = range(11)
# pmb_option / pmb_hierarchy_mode
OPT_HIERARCHY, OPT_FUSED_TAIL, OPT_TRACE = 0, 1, 2
HIERARCHY_AUTO, HIERARCHY_CHAINS, HIERARCHY_WAVEFRONT = 0, 1, 2
# pmb_cull_mode
CULL_AABB, CULL_SPHERE, CULL_FILTER_ONLY = 0, 1, 2
# pmb_kernel_id
K_FRAME, K_RESERVED, K_EXPAND, K_CULL, K_COMPACT, K_AUX = range(6)
KERNEL_NAMES = ["frame", "reserved", "expand", "cull", "compact", "aux", "anim", "draw"]

_ARRAY_SPEC = {  # id -> (dtype, trailing shape)
    PARENTS: (np.uint32, ()),
    LOCAL_MATRICES: (np.float32, (16,)),
    WORLD_MATRICES: (np.float32, (16,)),
    ENTITIES: (np.uint64, ()),
    STATE_FLAGS: (np.uint64, ()),
    BOUNDING_VOLUMES: (np.float32, (13,)),
    POS_EXTENT: (np.float32, (8,)),
    HAS_RB_MATRIX: (np.uint8, ()),
    TRANSFORMS: (np.float32, (10,)),  # translation xyz, rotation xyzw, scale xyz
    CBUFFER: (np.uint32, ()),
    DRAW_CALLS: (np.float32, (40,)),  # world 0:16, v1 16:20, v2 20:24, world_inv_transpose 24:40
}

# every symbol include/pmtech_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "pmb_scene_create", "pmb_scene_resize", "pmb_scene_destroy", "pmb_scene_capacity", "pmb_scene_device",
    "pmb_set_stream", "pmb_set_option", "pmb_download_trace", "pmb_synchronize", "pmb_upload", "pmb_download", "pmb_device_ptr", "pmb_device_bounds",
    "pmb_host_alloc", "pmb_host_free", "pmb_update_scene", "pmb_get_renderable_extents", "pmb_cull_views", "pmb_cull_views_static",
    "pmb_frame", "pmb_get_counts", "pmb_get_list", "pmb_get_lists", "pmb_get_lists_sb", "pmb_device_list", "pmb_device_counts",
    "pmb_copy_segments", "pmb_pack_lists", "pmb_pack_row", "pmb_unpack_rows",
    "pmb_publish_counts", "pmb_push_row", "pmb_update_draw_calls",
    "pmb_instance_blocks", "pmb_device_draw_calls", "pmb_download_instance_block",
    "pmb_exchange_create", "pmb_exchange_destroy", "pmb_exchange_ipc_handle", "pmb_exchange_connect_ipc",
    "pmb_exchange_connect_local", "pmb_exchange_launch", "pmb_exchange_flush", "pmb_exchange_finish", "pmb_exchange_get_list", "pmb_exchange_stream", "pmb_exchange_launches", "pmb_exchange_use_polling_kernels", "pmb_exchange_debug_header",
    "pmb_group_create", "pmb_group_destroy", "pmb_group_size", "pmb_group_scene", "pmb_group_exchange",
    "pmb_group_exchange_setup", "pmb_group_frame", "pmb_group_finish",
    "pmb_apply_root_motion", "pmb_skins_set", "pmb_skin_palettes", "pmb_download_palettes", "pmb_device_palettes",
    "pmb_frame_host", "pmb_set_profiling", "pmb_kernel_stats", "pmb_reset_kernel_stats", "pmb_total_launches",
    "pmb_camera_create_perspective", "pmb_camera_create_orthographic", "pmb_camera_create_cubemap",
    "pmb_camera_set_cubemap_face", "pmb_camera_look_at", "pmb_camera_update_frustum",
    "pmb_camera_update_shadow_frustum", "pmb_pms_parse", "pmb_load_pms",
    "pmb_mark_transforms_dirty", "pmb_reserve_views", "pmb_cull_shadow_views", "pmb_cull_spot_shadow_views", "pmb_camera_spot_shadow", "pmb_download_shadow_cameras", "pmb_device_shadow_cameras", "pmb_last_error", "pmb_version", "pmb_device_count",
]

# pmb_pms_component
PMS_COMPONENTS = ["entities", "state_flags", "parents", "transforms", "local_matrices", "world_matrices", "bounding_volumes",
                  "cbuffer", "draw_call_data", "pos_extent"]
PMS_SPEC = [(np.uint64, ()), (np.uint64, ()), (np.uint32, ()), (np.float32, (10,)), (np.float32, (16,)), (np.float32, (16,)),
            (np.float32, (13,)), (np.uint32, ()), (np.float32, (40,)), (np.float32, (8,))]


MASTER_INSTANCE_DTYPE = np.dtype([("num_instances", np.uint32), ("instance_buffer", np.uint32), ("instance_stride", np.uint32)])
INSTANCE_BLOCK_DTYPE = np.dtype([("master", np.uint32), ("instance_buffer", np.uint32), ("byte_offset", np.uint64), ("bytes", np.uint64)])
MAX_VIEWS, MAX_PEERS, IPC_HANDLE_BYTES = 16, 16, 64
XCHG_ROOT, XCHG_ALL, XCHG_COUNTS = 0, 1, 2


class ExchangeResult(C.Structure):
    """pmb_exchange_result (include/pmtech_b200.h)"""
    _fields_ = [("num_views", C.c_uint32), ("world", C.c_uint32), ("overflow", C.c_uint32), ("epoch", C.c_uint32),
                ("view_offset", C.c_uint32 * MAX_VIEWS), ("view_total", C.c_uint32 * MAX_VIEWS),
                ("counts", (C.c_uint32 * MAX_VIEWS) * MAX_PEERS), ("renderable_extents", C.c_float * 6),
                ("device_ms", C.c_float), ("lists_device", C.c_void_p)]


class PmsInfo(C.Structure):
    """pmb_pms_info (include/pmtech_b200.h)"""
    _fields_ = [("version", C.c_int32), ("num_nodes", C.c_uint32), ("num_components", C.c_int32),
                ("num_base_components", C.c_int32), ("num_lookup_strings", C.c_int32), ("num_extensions", C.c_int32),
                ("view_flags", C.c_uint32), ("selected_index", C.c_int32), ("num_cameras", C.c_uint32), ("reserved", C.c_uint32),
                ("sizes_offset", C.c_uint64), ("strings_offset", C.c_uint64), ("cameras_offset", C.c_uint64),
                ("data_offset", C.c_uint64), ("names_offset", C.c_uint64), ("component_offset", C.c_uint64 * len(PMS_COMPONENTS))]


SKIN_DTYPE = np.dtype([("joints_offset", np.uint32), ("num_joints", np.uint32), ("bind_offset", np.uint32),
                       ("palette_offset", np.uint32)])


class PmbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pmtech_b200 error {code}: {msg}")
        self.code = code


class HostArrays(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "parents", "local_matrices", "entities", "state_flags", "bounding_volumes", "has_rb_matrix", "transforms",
        "world_matrices_out", "bounding_volumes_out", "pos_extent_out", "renderable_extents_out")]


class Segment(C.Structure):
    _fields_ = [("src_offset", C.c_uint64), ("dst_offset", C.c_uint64), ("count", C.c_uint32), ("add", C.c_uint32)]


SEGMENT_DTYPE = np.dtype([("src_offset", "<u8"), ("dst_offset", "<u8"), ("count", "<u4"), ("add", "<u4")])


class CameraStruct(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("view", C.c_float * 16), ("proj", C.c_float * 16),
                ("camera_frustum", C.c_float * FRUSTUM_FLOATS)]


_lib = None


def load():
    """dlopen the library and declare the prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RuntimeError(f"{SO_PATH} is missing: run `python -m pmtech_b200.build` (nvcc, sm_100a). "
                           "There is no CPU fallback for this path.")
    L = C.CDLL(SO_PATH)
    vp, u32, i32 = C.c_void_p, C.c_uint32, C.c_int
    L.pmb_scene_create.argtypes = [C.POINTER(vp), u32, i32]
    L.pmb_scene_resize.argtypes = [vp, u32]
    L.pmb_scene_destroy.argtypes = [vp]
    L.pmb_scene_destroy.restype = None
    L.pmb_scene_capacity.argtypes = [vp]
    L.pmb_scene_capacity.restype = u32
    L.pmb_scene_device.argtypes = [vp]
    L.pmb_set_stream.argtypes = [vp, vp]
    L.pmb_synchronize.argtypes = [vp]
    L.pmb_set_option.argtypes = [vp, i32, i32]
    L.pmb_download_trace.argtypes = [vp, vp, u32]
    L.pmb_upload.argtypes = [vp, i32, vp, u32, u32]
    L.pmb_download.argtypes = [vp, i32, vp, u32, u32]
    L.pmb_device_ptr.argtypes = [vp, i32]
    L.pmb_device_ptr.restype = vp
    L.pmb_device_bounds.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.pmb_host_alloc.argtypes = [C.c_size_t]
    L.pmb_host_alloc.restype = vp
    L.pmb_host_free.argtypes = [vp]
    L.pmb_host_free.restype = None
    L.pmb_update_scene.argtypes = [vp, u32]
    L.pmb_get_renderable_extents.argtypes = [vp, vp]
    L.pmb_cull_views.argtypes = [vp, u32, vp, u32, i32]
    L.pmb_cull_views_static.argtypes = [vp, vp, vp, u32, vp, u32, i32, vp, C.c_uint64, vp, vp]
    L.pmb_cull_views_static.restype = i32
    L.pmb_frame.argtypes = [vp, u32, vp, u32, i32]
    L.pmb_get_counts.argtypes = [vp, vp, u32]
    L.pmb_get_list.argtypes = [vp, u32, vp, u32]
    L.pmb_get_list.restype = C.c_int64
    L.pmb_get_lists.argtypes = [vp, C.POINTER(vp), vp, vp, u32]
    L.pmb_get_lists_sb.argtypes = [vp, C.POINTER(vp), u32]
    L.pmb_device_list.argtypes = [vp, u32]
    L.pmb_device_list.restype = vp
    L.pmb_device_counts.argtypes = [vp]
    L.pmb_device_counts.restype = vp
    L.pmb_copy_segments.argtypes = [vp, vp, vp, vp, u32]
    L.pmb_pack_lists.argtypes = [vp, vp, u32, u32, vp]
    L.pmb_pack_row.argtypes = [vp, u32, u32, vp, u32]
    L.pmb_publish_counts.argtypes = [vp, vp, u32, u32, u32, vp, vp]
    L.pmb_push_row.argtypes = [vp, vp, u32, u32, u32, u32, vp, vp, C.c_uint64, vp, vp]
    L.pmb_update_draw_calls.argtypes = [vp, u32, C.c_float]
    L.pmb_instance_blocks.argtypes = [vp, vp, u32, vp, u32]
    L.pmb_instance_blocks.restype = C.c_int64
    L.pmb_device_draw_calls.argtypes = [vp]
    L.pmb_device_draw_calls.restype = vp
    L.pmb_download_instance_block.argtypes = [vp, vp, vp]
    u64 = C.c_uint64
    L.pmb_exchange_create.argtypes = [C.POINTER(vp), vp, u32, u32, u32, u32, u64, u64, i32, u32]
    L.pmb_exchange_destroy.argtypes = [vp]
    L.pmb_exchange_destroy.restype = None
    L.pmb_exchange_ipc_handle.argtypes = [vp, vp]
    L.pmb_exchange_connect_ipc.argtypes = [vp, vp]
    L.pmb_exchange_connect_local.argtypes = [C.POINTER(vp), u32]
    L.pmb_exchange_launch.argtypes = [vp]
    L.pmb_exchange_flush.argtypes = [vp]
    L.pmb_exchange_finish.argtypes = [vp, C.POINTER(ExchangeResult)]
    L.pmb_exchange_get_list.argtypes = [vp, u32, vp, u32]
    L.pmb_exchange_get_list.restype = C.c_int64
    L.pmb_exchange_use_polling_kernels.argtypes = [i32]
    L.pmb_exchange_use_polling_kernels.restype = None
    L.pmb_exchange_debug_header.argtypes = [vp, vp, u32]
    L.pmb_exchange_stream.argtypes = [vp]
    L.pmb_exchange_stream.restype = vp
    L.pmb_exchange_launches.argtypes = [vp]
    L.pmb_exchange_launches.restype = u64
    L.pmb_group_create.argtypes = [C.POINTER(vp), vp, u32, vp]
    L.pmb_group_destroy.argtypes = [vp]
    L.pmb_group_destroy.restype = None
    L.pmb_group_size.argtypes = [vp]
    L.pmb_group_size.restype = u32
    L.pmb_group_scene.argtypes = [vp, u32]
    L.pmb_group_scene.restype = vp
    L.pmb_group_exchange.argtypes = [vp, u32]
    L.pmb_group_exchange.restype = vp
    L.pmb_group_exchange_setup.argtypes = [vp, u32, vp, u64, u64, i32, u32]
    L.pmb_group_frame.argtypes = [vp, vp, vp, u32, i32]
    L.pmb_group_finish.argtypes = [vp, u32, C.POINTER(ExchangeResult)]
    L.pmb_apply_root_motion.argtypes = [vp, vp, vp, vp, u32]
    L.pmb_skins_set.argtypes = [vp, vp, u32, vp, u32, u32]
    L.pmb_skin_palettes.argtypes = [vp]
    L.pmb_download_palettes.argtypes = [vp, vp, u32, u32]
    L.pmb_device_palettes.argtypes = [vp]
    L.pmb_device_palettes.restype = vp
    L.pmb_unpack_rows.argtypes = [vp, vp, u32, u32, u32, vp, C.c_uint64, vp, vp]
    L.pmb_frame_host.argtypes = [vp, u32, C.POINTER(HostArrays), vp, u32, i32, C.POINTER(vp), vp]
    L.pmb_set_profiling.argtypes = [vp, i32]
    L.pmb_kernel_stats.argtypes = [vp, i32, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    L.pmb_reset_kernel_stats.argtypes = [vp]
    L.pmb_total_launches.argtypes = [vp]
    L.pmb_total_launches.restype = C.c_uint64
    cam = C.POINTER(CameraStruct)
    f = C.c_float
    L.pmb_camera_create_perspective.argtypes = [cam, f, f, f, f]
    L.pmb_camera_create_orthographic.argtypes = [cam, f, f, f, f, f, f]
    L.pmb_camera_create_cubemap.argtypes = [cam, f, f]
    L.pmb_camera_set_cubemap_face.argtypes = [cam, u32]
    L.pmb_camera_look_at.argtypes = [cam, vp, vp, vp]
    L.pmb_camera_update_frustum.argtypes = [cam]
    L.pmb_camera_update_shadow_frustum.argtypes = [cam, vp, vp, vp]
    L.pmb_camera_spot_shadow.argtypes = [cam, vp]
    L.pmb_camera_spot_shadow.restype = None
    for fn in ("pmb_camera_create_perspective", "pmb_camera_create_orthographic", "pmb_camera_create_cubemap",
               "pmb_camera_set_cubemap_face", "pmb_camera_look_at", "pmb_camera_update_frustum",
               "pmb_camera_update_shadow_frustum"):
        getattr(L, fn).restype = None
    L.pmb_mark_transforms_dirty.argtypes = [vp, u32, u32]
    L.pmb_reserve_views.argtypes = [vp, u32]
    L.pmb_cull_shadow_views.argtypes = [vp, u32, vp, u32, vp, i32, u32]
    L.pmb_cull_spot_shadow_views.argtypes = [vp, u32, vp, u32, i32, u32]
    L.pmb_download_shadow_cameras.argtypes = [vp, vp, u32]
    L.pmb_device_shadow_cameras.argtypes = [vp]
    L.pmb_device_shadow_cameras.restype = vp
    L.pmb_pms_parse.argtypes = [vp, C.c_uint64, C.POINTER(PmsInfo)]
    L.pmb_load_pms.argtypes = [vp, vp, C.c_uint64, u32, C.POINTER(PmsInfo)]
    L.pmb_last_error.restype = C.c_char_p
    L.pmb_version.restype = C.c_char_p
    _lib = L
    return L


def _check(rc):
    if rc < 0:
        raise PmbError(rc, load().pmb_last_error().decode())
    return rc


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def device_count() -> int:
    return load().pmb_device_count()


def pinned_empty(shape, dtype):
    """numpy array over cudaHostAlloc'ed (page-locked) memory; freed with pinned_free()."""
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape)) * dtype.itemsize
    p = load().pmb_host_alloc(max(nbytes, 1))
    if not p:
        raise PmbError(-4, load().pmb_last_error().decode())
    buf = (C.c_char * max(nbytes, 1)).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    arr.flags.writeable = True
    _PINNED[arr.ctypes.data] = p
    return arr


_PINNED = {}


def pinned_free(arr):
    p = _PINNED.pop(arr.ctypes.data, None)
    if p:
        load().pmb_host_free(p)
