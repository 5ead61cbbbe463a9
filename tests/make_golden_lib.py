"""Scenes behind the committed fixtures of tests/golden/ (shared by the generator script and the
tests that replay them).  Small on purpose: a fixture is a few hundred KB."""
import numpy as np


def scene_for(name, camera_cls):
    from pmtech_b200 import synth
    if name == "mixed_3000":
        return synth.mixed_scene(n=3000, seed=42), synth.mixed_views(camera_cls), "aabb"
    if name == "torus_views":
        sc = synth.torus_scene()
        keep = 4104  # lights, floor, walls, master + the first 16 rows of boxes: closed under parent links
        sc = synth.slice_scene(sc, 0, keep)
        return sc, synth.torus_views(camera_cls), "aabb"
    if name == "rigs_40":
        return synth.rig_scene(40, seed=4), synth.mixed_views(camera_cls), "sphere"
    raise KeyError(name)


NAMES = ["mixed_3000", "torus_views", "rigs_40"]


def stage0_case(camera_cls, characters=30, seed=8):
    """The animated-rig frame behind tests/golden/rigs_stage0.npz: every joint dirtied (e_cmp::transform),
    one root-motion delta per rig, sphere cull, skin palettes and draw-call records (a third of the
    entities without a cbuffer handle, the meshes flagged skinned)."""
    from pmtech_b200 import synth
    sc = synth.rig_scene(characters, seed=seed)
    n = sc["num_entities"]
    sc["entities"] = sc["entities"] | np.uint64(1 << 12)
    mesh = (np.arange(n) % synth.RIG_ENTITIES) == 1
    sc["entities"][mesh] |= np.uint64(1 << 6)  # skinned: draw call carries identity (ecs_scene.cpp:1985)
    cb = (np.arange(n, dtype=np.uint32) % 3) * np.uint32(7)  # 0 (null) for every third entity
    cb[5::11] = 0xFFFFFFFF
    rng = np.random.default_rng(seed)
    dc0 = rng.standard_normal((n, 40)).astype(np.float32)
    return sc, synth.mixed_views(camera_cls), cb, dc0, np.float32(4321.5)


def run_stage0_oracle(orc, case):
    sc, views, cb, dc0, time_ms = case
    rm = sc["root_motion"]
    o = orc.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                        transforms=sc["transforms"])
    o.apply_root_motion(rm["targets"], rm["rotations"], rm["deltas"])
    lists, _ = o.frame(views, mode="sphere")
    pal = o.skin_palettes(sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])
    dc = o.update_draw_calls(cb, dc0.copy(), float(time_ms))
    return dict(transforms=o.transforms, local=o.local, world=o.world, bv=o.bv, pos_extent=o.pos_extent,
                entities=o.entities, palettes=pal, draw_calls=dc, lists=lists)


def run_stage0_device(case):
    from pmtech_b200 import capi
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc, views, cb, dc0, time_ms = case
    n, rm = sc["num_entities"], sc["root_motion"]
    with DeviceScene(n, 0) as dev:
        upload_scene(dev, sc)
        dev.upload(capi.CBUFFER, cb)
        dev.upload(capi.DRAW_CALLS, dc0)
        dev.set_skins(sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])
        dev.apply_root_motion(rm["targets"], rm["rotations"], rm["deltas"])
        dev.frame(n, views, "sphere")
        dev.skin_palettes()
        dev.update_draw_calls(n, float(time_ms))
        return dict(transforms=dev.download(capi.TRANSFORMS, 0, n), local=dev.download(capi.LOCAL_MATRICES, 0, n),
                    world=dev.download(capi.WORLD_MATRICES, 0, n), bv=dev.download(capi.BOUNDING_VOLUMES, 0, n),
                    pos_extent=dev.download(capi.POS_EXTENT, 0, n), entities=dev.download(capi.ENTITIES, 0, n),
                    palettes=dev.download_palettes(), draw_calls=dev.download(capi.DRAW_CALLS, 0, n),
                    lists=dev.lists(len(views)))


STAGE0_KEYS = ("transforms", "local", "world", "bv", "pos_extent", "entities", "palettes", "draw_calls")
