#!/usr/bin/env python
"""Writes tests/golden/pmtech_sdf_scene.npz and pmtech_lucy_scene.npz from the scene files the reference
ships with its examples:

    /root/reference/examples/assets/scene/sdf.pms        (version 8, 23 nodes, 31 components)
    /root/reference/examples/assets/scene/lucy.obj.pms   (version 9,  1 node,  35 components)

These are OUTPUTS OF PMTECH ITSELF: save_scene dumps every component array as it stood after the editor's
last update_scene (ecs_scene.cpp:2162-2346), so next to the inputs of the path (parents, transforms, local
matrices, bounding-volume min/max, cbuffer handles) the file holds what pmtech computed from them — world
matrices, transformed extents and radii, parent-expanded boxes, draw-call records with the inverse
transposes.  They are the only reference-computed vectors this path has, and what the oracle is pinned
against (tests/test_pmtech_scene_files.py).  Only the path's component arrays, the node names
and the cameras are kept (a few KB); the reader is pmtech_b200/pms.py.  /root/reference exists only in the
build container, which is why the arrays are committed.

    python tests/golden/make_pms_fixture.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from pmtech_b200 import pms  # noqa: E402

SCENES = {"pmtech_sdf_scene": "sdf.pms", "pmtech_lucy_scene": "lucy.obj.pms"}
ROOT = "/root/reference/examples/assets/scene"


def main():
    for out_name, src in SCENES.items():
        p = pms.read_pms(os.path.join(ROOT, src))
        out = {k: v for k, v in p.components.items()}
        out["component_sizes"] = p.component_sizes
        out["version"] = np.int32(p.version)
        out["names"] = np.array(p.names)
        cams = np.array([[*c["pos"], *c["focus"], *c["rot"], c["fov"], c["aspect"], c["near"], c["far"], c["zoom"]]
                         for c in p.cameras], np.float32).reshape(-1, 13)
        out["cameras"] = cams
        out["file_bytes"] = np.int64(len(p.data))
        out["names_offset"] = np.int64(p.info.names_offset)
        path = os.path.join(HERE, out_name + ".npz")
        np.savez_compressed(path, **out)
        print(out_name, "version", p.version, p.num_nodes, "nodes", sorted(p.components), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
