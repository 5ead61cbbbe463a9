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
