#!/usr/bin/env python
"""Device time and achieved HBM GB/s of the secondary kernels at a size well past L2 (one JSON line each):
stage-0 bake (56 B read + 64 B written per baked entity), draw-call packing (64 + 8 + 4 B read, 136 B
written per live entity), skin palettes (64 B read + 64 B written per joint).

    python scripts/bench_kernels.py [--entities N] [--steps K]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--entities", type=int, default=1 << 23)
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    import torch
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    n = args.entities
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sc = synth.city_scene(n, seed=3, with_transforms=True)
    flag = np.uint64(1 << 12)
    ent_dirty = capi.pinned_empty((n,), np.uint64)
    ent_dirty[...] = sc["entities"] | flag

    def report(name, ms, bytes_per_unit, units, note):
        gbs = bytes_per_unit * units / (ms * 1e-3) / 1e9
        print(json.dumps({"kernel": name, "units": units, "ms": ms, "algorithmic_bytes_per_unit": bytes_per_unit,
                          "achieved_gbs": gbs, "peak_gbs": peak, "frac": gbs / peak, "note": note}), flush=True)

    with DeviceScene(n, 0) as dev:
        dev.set_stream(stream.cuda_stream)
        for aid, key in ((capi.PARENTS, "parents"), (capi.LOCAL_MATRICES, "local_matrices"), (capi.ENTITIES, "entities"),
                         (capi.STATE_FLAGS, "state_flags"), (capi.BOUNDING_VOLUMES, "bounding_volumes"), (capi.TRANSFORMS, "transforms")):
            dev.upload(aid, sc[key])
        dev.upload(capi.CBUFFER, np.ones(n, np.uint32))
        dev.synchronize()
        dev.update_scene(n)
        dev.synchronize()
        # ---- stage 0: every entity flagged
        dev.set_profiling(True)
        for it in range(args.steps + 3):
            if it == 3:
                dev.synchronize()
                dev.reset_kernel_stats()
            dev.upload(capi.ENTITIES, ent_dirty)
            dev.update_scene(n)
        dev.synchronize()
        ms, cnt = dev.kernel_stats()["anim"]
        report("bake_kernel", ms / cnt, 120, n, "all entities flagged e_cmp::transform: 40 B transform + 16 B flags in, 64 B local out")
        # ---- draw calls
        dev.reset_kernel_stats()
        for _ in range(args.steps):
            dev.update_draw_calls(n, 1.0)
        dev.synchronize()
        ms, cnt = dev.kernel_stats()["draw"]
        report("draw_call_kernel", ms / cnt, 64 + 8 + 4 + 136, n, "every entity live: world + flags + handle in, world + v1.xy + inverse transpose out; "
               "records never uploaded by the host (v1.zw / v2 known to be zero)")
        dev.upload(capi.DRAW_CALLS, np.zeros((1, 40), np.float32))  # from now on v1.zw / v2 are the user's and are read back
        dev.reset_kernel_stats()
        for _ in range(args.steps):
            dev.update_draw_calls(n, 1.0)
        dev.synchronize()
        ms, cnt = dev.kernel_stats()["draw"]
        report("draw_call_kernel<user data>", ms / cnt, 64 + 8 + 4 + 32 + 136, n, "the same with v1 / v2 (32 B of every 160 B record) read back first")
        # ---- skin palettes: n / 64 meshes of 64 joints, bind matrices shared by all meshes
        k = n // 64
        sk = np.zeros(k, capi.SKIN_DTYPE)
        sk["joints_offset"] = np.arange(k, dtype=np.uint32) * 64
        sk["num_joints"] = 64
        sk["palette_offset"] = np.arange(k, dtype=np.uint32) * 64
        dev.set_skins(sk, synth.trs_matrices(np.zeros((64, 3), np.float32), synth.random_unit_quats(np.random.default_rng(1), 64),
                                             np.ones((64, 3), np.float32)), k * 64)
        dev.reset_kernel_stats()
        for _ in range(args.steps):
            dev.skin_palettes()
        dev.synchronize()
        ms, cnt = dev.kernel_stats()["anim"]
        report("skin_kernel", ms / cnt, 128, k * 64, "64-joint meshes, one shared skin: 64 B joint world in, 64 B palette matrix out")
        dev.set_profiling(False)


if __name__ == "__main__":
    main()
