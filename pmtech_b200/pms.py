"""`.pms` scene files for the scene-update path.

Host-side mirror of put::ecs::load_scene / save_scene (core/put/source/ecs/ecs_scene.cpp:2350-2535,
:2162-2346) restricted to what the path consumes: the walk over the image is done by the library
(`pmb_pms_parse`, pm_pms.cpp); this module wraps it, hands out numpy views of the component arrays in
the reference's layouts, reads the lookup strings / cameras / node names, and writes images of the
same layout for scenes built in code.  Renderer resources (geometry, materials, animations; the
specialisation blocks after the node names) are the engine's business and are not touched.
"""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np

from . import capi

HEADER_BYTES = 256  # scene_header: 64 x s32 (ecs_scene.cpp:2031-2044)
# element sizes of ecs_scene.h:471-513 as current pmtech writes them (index = component number);
# 0 = a component this module never needs to size itself
BASE_COMPONENT_SIZES_V8 = (8, 8, 4, 4, 4, 16, 16, 16, 4, 40, 64, 64, 64, 64, 52, 52, 4, 12, 48, 16, 160, 32, 4, 160, 24, 16,
                           256, 352, 8, 192, 4)
_PATH_INDEX = {"entities": 0, "state_flags": 1, "parents": 8, "transforms": 9, "local_matrices": 10, "world_matrices": 11,
               "bounding_volumes": 14, "cbuffer": 22, "draw_call_data": 23, "pos_extent": 38}


class PmsScene:
    """A parsed image: `header` fields, `component_sizes`, `lookup_strings` {hash: str}, `cameras`, `names`,
    and one numpy view per path component found (`.components[name]`, absent when the file stores another
    element size — the reference skips those too, ecs_scene.cpp:2506-2524)."""

    def __init__(self, data: bytes):
        self.data = bytes(data)
        info = capi.PmsInfo()
        buf = np.frombuffer(self.data, np.uint8)
        capi._check(capi.load().pmb_pms_parse(capi._ptr(buf), len(self.data), C.byref(info)))
        self.info = info
        self.version, self.num_nodes = info.version, info.num_nodes
        n = self.num_nodes
        self.component_sizes = np.frombuffer(self.data, np.uint32, info.num_components, info.sizes_offset).copy()
        self.components = {}
        for k, name in enumerate(capi.PMS_COMPONENTS):
            off = info.component_offset[k]
            if off:
                dt, tail = capi.PMS_SPEC[k]
                count = n * int(np.prod(tail, dtype=np.int64))
                self.components[name] = np.frombuffer(self.data, dt, count, off).reshape((n,) + tail)
        # lookup strings: u32 length, bytes, hash_id (write_lookup_string / read_parsable_string)
        self.lookup_strings = {}
        off = info.strings_offset
        for _ in range(info.num_lookup_strings):
            (ln,) = struct.unpack_from("<I", self.data, off)
            s = self.data[off + 4:off + 4 + ln].decode("latin1")
            (hid,) = struct.unpack_from("<I", self.data, off + 4 + ln)
            self.lookup_strings[hid] = s
            off += 8 + ln
        self.cameras = []
        off = info.cameras_offset
        for _ in range(info.num_cameras):  # :2449-2457
            v = struct.unpack_from("<I13f", self.data, off)
            self.cameras.append(dict(id=v[0], pos=v[1:4], focus=v[4:7], rot=v[7:9], fov=v[9], aspect=v[10], near=v[11],
                                     far=v[12], zoom=v[13]))
            off += 56
        # per node: name, geometry name, material name as lookup hashes (:2534-2542)
        self.names = []
        off = info.names_offset
        if off + 12 * n <= len(self.data):
            ids = np.frombuffer(self.data, np.uint32, 3 * n, off).reshape(n, 3)
            self.names = [self.lookup_strings.get(int(i), "") for i in ids[:, 0]]

    def host_arrays(self, zero_offset=0):
        """The path's inputs as fresh arrays ready for upload (parents += zero_offset as a merge does, :2530)."""
        out = {k: v.copy() for k, v in self.components.items()}
        if "parents" in out and zero_offset:
            out["parents"] = out["parents"] + np.uint32(zero_offset)
        out["num_entities"] = self.num_nodes
        return out


def read_pms(path) -> PmsScene:
    with open(path, "rb") as f:
        return PmsScene(f.read())


def load_into(dev, data: bytes, zero_offset=0) -> capi.PmsInfo:
    """load_scene for the path: every component found goes straight to the device mirrors of `dev`
    (a DeviceScene) at [zero_offset, zero_offset + num_nodes)."""
    info = capi.PmsInfo()
    buf = np.frombuffer(data, np.uint8)
    capi._check(capi.load().pmb_load_pms(dev._h, capi._ptr(buf), len(data), int(zero_offset), C.byref(info)))
    dev.synchronize()
    return info


def write_pms(arrays: dict, version=10, lookup_strings=(), cameras=(), view_flags=0, selected_index=0) -> bytes:
    """save_scene's layout (:2162-2346) for the components in `arrays` (names of capi.PMS_COMPONENTS); every
    other component is written with element size 0, which readers skip.  `version` < 10 drops pos_extent
    (component 38 arrived with version 10, ecs_scene.h:509)."""
    n = int(arrays["num_entities"]) if "num_entities" in arrays else int(np.asarray(arrays["entities"]).shape[0])
    ncomp = 39 if version >= 10 else len(BASE_COMPONENT_SIZES_V8)
    sizes = np.zeros(ncomp, np.uint32)
    blobs = {}
    for k, name in enumerate(capi.PMS_COMPONENTS):
        if name in arrays and _PATH_INDEX[name] < ncomp:
            dt, tail = capi.PMS_SPEC[k]
            a = np.ascontiguousarray(np.asarray(arrays[name])[:n], dt).reshape((n,) + tail)
            blobs[_PATH_INDEX[name]] = a.tobytes()
            sizes[_PATH_INDEX[name]] = a.dtype.itemsize * int(np.prod(tail, dtype=np.int64))
    header = np.zeros(64, np.int32)
    header[0], header[1], header[2], header[3] = HEADER_BYTES, version, n, ncomp
    header[4], header[5], header[6] = len(lookup_strings), 0, ncomp
    header[32], header[33] = view_flags, selected_index
    out = [header.tobytes(), sizes.tobytes()]
    for s, hid in lookup_strings:
        b = s.encode("latin1")
        out.append(struct.pack("<I", len(b)) + b + struct.pack("<I", hid))
    out.append(struct.pack("<I", len(cameras)))
    for c in cameras:
        out.append(struct.pack("<I13f", c["id"], *c["pos"], *c["focus"], *c["rot"], c["fov"], c["aspect"], c["near"], c["far"],
                               c["zoom"]))
    for i in range(ncomp):
        out.append(blobs.get(i, b""))
    out.append(np.zeros(3 * n, np.uint32).tobytes())  # name / geometry / material lookups: the empty string
    return b"".join(out)
