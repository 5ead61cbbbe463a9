#!/usr/bin/env python
"""Device timings of the BASELINE configs other than the headline one (which bench.py measures):
config 1 (64k torus, 1 camera), config 2 (1M entities, depth 16, update only), config 4 (10k rigs of
64 bones: root motion + stage-0 bake + propagation + skin palettes).  One JSON line per config.

    python scripts/bench_configs.py [--steps K]
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
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--only", type=int, nargs="*", default=[1, 2, 4], help="which BASELINE configs to run")
    ap.add_argument("--no-tail", action="store_true", help="PMB_OPT_FUSED_TAIL = 0: separate launches")
    ap.add_argument("--force-tail", action="store_true", help="PMB_OPT_FUSED_TAIL = 2: one launch whenever the scene is small enough")
    ap.add_argument("--hierarchy", default="auto", choices=["auto", "chains", "wavefront"])
    args = ap.parse_args()
    import torch
    from pmtech_b200 import capi, synth
    from pmtech_b200.ecs import Camera
    from pmtech_b200.scene import DeviceScene
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)

    def run(name, sc, views, mode, pre=None, post=None, bytes_per_entity=232):
        n = sc["num_entities"]
        with DeviceScene(n, 0) as dev:
            dev.set_stream(stream.cuda_stream)
            dev.set_option(capi.OPT_FUSED_TAIL, 0 if args.no_tail else (2 if args.force_tail else 1))
            dev.set_option(capi.OPT_HIERARCHY, {"auto": 0, "chains": 1, "wavefront": 2}[args.hierarchy])
            keep = [dev.upload(capi.PARENTS, sc["parents"]), dev.upload(capi.LOCAL_MATRICES, sc["local_matrices"]),
                    dev.upload(capi.ENTITIES, sc["entities"]), dev.upload(capi.STATE_FLAGS, sc["state_flags"]),
                    dev.upload(capi.BOUNDING_VOLUMES, sc["bounding_volumes"])]
            if "transforms" in sc:
                keep.append(dev.upload(capi.TRANSFORMS, sc["transforms"]))
            dev.synchronize()

            def step():
                if pre:
                    pre(dev)
                if views is not None and len(views):
                    dev.frame(n, views, mode)
                else:
                    dev.update_scene(n)
                if post:
                    post(dev)
            for _ in range(5):
                step()
            torch.cuda.synchronize()
            dev.reset_kernel_stats()
            dev.set_profiling(True)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.steps):
                step()
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            stats = {k: v[0] / args.steps for k, v in dev.kernel_stats().items() if v[1]}
            dev.set_profiling(False)
        print(json.dumps({"config": name, "entities": n, "views": len(views) if views is not None else 0, "ms_per_step": ms,
                          "entities_per_s": n / (ms * 1e-3), "step_gbs_232": bytes_per_entity * n / (ms * 1e-3) / 1e9,
                          "kernel_ms_by_group": stats}), flush=True)

    if 1 in args.only:
        run("config1: 64k torus, 1 camera", synth.torus_scene(seed=1), synth.torus_views(Camera), "aabb")
    if 2 in args.only:
        run("config2: 1M entities, depth 16, update only", synth.deep_scene(), None, "aabb")
    if 4 not in args.only:
        return
    rig = synth.rig_scene(10_000, seed=4)
    rm = rig["root_motion"]
    n_rig = rig["num_entities"]
    rm_t, rm_q, rm_d = (np.ascontiguousarray(rm["targets"], np.uint32), np.ascontiguousarray(rm["rotations"], np.float32),
                        np.ascontiguousarray(rm["deltas"], np.float32))

    def pre(dev):  # the animation dirtied every joint (e_cmp::transform set for the range on the device) and moved every rig root
        dev.mark_transforms_dirty(0, n_rig)
        dev.apply_root_motion(rm_t, rm_q, rm_d, sync=False)

    def setup_skins(dev):
        if not getattr(dev, "_skins_set", False):
            dev.set_skins(rig["skins"], rig["bind_matrices"], rig["num_palette_matrices"])
            dev._skins_set = True
        dev.skin_palettes()
    run("config4: 10k rigs x 64 bones, every joint dirtied: root motion + bake + propagation + 6-view sphere cull + palettes",
        rig, synth.mixed_views(Camera), "sphere", pre=pre, post=setup_skins)


if __name__ == "__main__":
    main()
