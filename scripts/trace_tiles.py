#!/usr/bin/env python
"""Timeline of the wavefront form's tiles (PMB_OPT_TRACE): when each 64-entity tile started loop A, when it was through,
how many rounds its plan has and how many of them need nothing from other tiles.  python scripts/trace_tiles.py [2|4] > gpurun_out/trace.json"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    sc = synth.deep_scene() if cfg == 2 else synth.rig_scene(10_000, seed=4)
    n = sc["num_entities"]
    tiles = (n + 63) // 64
    with DeviceScene(n, 0) as dev:
        dev.set_option(capi.OPT_TRACE, 1)
        for a, k in ((capi.PARENTS, "parents"), (capi.LOCAL_MATRICES, "local_matrices"), (capi.ENTITIES, "entities"),
                     (capi.STATE_FLAGS, "state_flags"), (capi.BOUNDING_VOLUMES, "bounding_volumes")):
            dev.upload(a, sc[k])
        for _ in range(3):
            dev.update_scene(n)
        dev.synchronize()
        tr = dev.download_trace(tiles)
    t0 = tr[:, 0].astype(np.int64)
    t1 = tr[:, 1].astype(np.int64)
    org = t0.min()
    start, end = (t0 - org) / 1e3, (t1 - org) / 1e3
    batches = (tr[:, 2] >> np.uint64(32)).astype(np.int64)
    polls = (tr[:, 3] & np.uint64(0xFFFFFFFF)).astype(np.int64)  # rounds of phase 0 (local)
    dur = end - start
    out = {"config": cfg, "tiles": int(tiles), "span_us": float(end.max()),
           "loop_a_us": {"median": float(np.median(dur)), "p90": float(np.percentile(dur, 90)), "max": float(dur.max())},
           "rounds": {"median": float(np.median(batches)), "max": int(batches.max())},
           "local_rounds": {"median": float(np.median(polls)), "p90": float(np.percentile(polls, 90)), "max": int(polls.max())},
           "start_us_by_tile_decile": [float(start[int(i * (tiles - 1) / 10)]) for i in range(11)],
           "end_us_by_tile_decile": [float(end[int(i * (tiles - 1) / 10)]) for i in range(11)],
           "first_64_tiles": [[round(float(start[i]), 2), round(float(end[i]), 2), int(batches[i]), int(polls[i])] for i in range(64)]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
