#!/usr/bin/env python
"""Writes tests/golden/*.npz: inputs' views and the oracle's outputs for the scenes of
tests/make_golden_lib.py.

These vectors come from oracle/pm_oracle.c (the CPU restatement), cross-checked bit for bit against
tests/numpy_ref.py before they are written; they are NOT outputs of the reference itself, which does
not compile here (third_party/maths is absent from /root/reference).  They freeze the
oracle's arithmetic so that the CUDA path, the oracle and the numpy restatement cannot drift apart
silently, and they travel to the GPU box.  The reference-computed vectors that pin the oracle itself
are the scene-file fixtures of make_pms_fixture.py.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import make_golden_lib as mg  # noqa: E402
import numpy_ref  # noqa: E402
from helpers import run_oracle  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def main():
    orc.build()
    for name in mg.NAMES:
        sc, views, mode = mg.scene_for(name, orc.OracleCamera)
        got = run_oracle(orc, sc, views, mode=mode)
        ref = numpy_ref.frame(sc, views, mode)
        for k in ("world", "bv", "pos_extent", "extents"):
            assert np.array_equal(np.asarray(got[k]).view(np.uint32), np.asarray(ref[k]).view(np.uint32)), (name, k)
        for v in range(len(views)):
            assert np.array_equal(got["lists"][v], ref["lists"][v]), (name, v)
        out = dict(world=got["world"], bv=got["bv"], pos_extent=got["pos_extent"], extents=got["extents"],
                   entities=got["entities"], state_flags=got["state_flags"], views=views)
        for v, l in enumerate(got["lists"]):
            out[f"list{v}"] = l
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, sc["num_entities"], "entities", [len(l) for l in got["lists"]], os.path.getsize(path), "bytes")
    stage0()


def stage0():
    """rigs_stage0.npz: root motion + stage-0 bake + frame + palettes + draw calls, cross-checked piece by
    piece against the numpy restatement before it is written."""
    case = mg.stage0_case(orc.OracleCamera)
    sc, views, cb, dc0, time_ms = case
    got = mg.run_stage0_oracle(orc, case)
    rm = sc["root_motion"]
    t_ref, e_ref = numpy_ref.apply_root_motion(sc, rm["targets"], rm["rotations"], rm["deltas"])
    local_ref, t_baked = numpy_ref.bake_transforms(dict(sc, transforms=t_ref, entities=e_ref))
    assert np.array_equal(got["transforms"].view(np.uint32), t_baked.view(np.uint32))  # rotations normalised in place
    assert np.array_equal(got["local"].view(np.uint32), local_ref.view(np.uint32))
    pal_ref = numpy_ref.skin_palettes(got["world"], sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])
    assert np.array_equal(got["palettes"].view(np.uint32), pal_ref.view(np.uint32))
    dc_ref = numpy_ref.update_draw_calls(got["world"], got["entities"], cb, dc0, time_ms)
    assert np.array_equal(got["draw_calls"].view(np.uint32), dc_ref.view(np.uint32))
    out = {k: got[k] for k in mg.STAGE0_KEYS}
    for v, l in enumerate(got["lists"]):
        out[f"list{v}"] = l
    path = os.path.join(HERE, "rigs_stage0.npz")
    np.savez_compressed(path, **out)
    print("rigs_stage0", sc["num_entities"], "entities", [len(l) for l in got["lists"]], os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
