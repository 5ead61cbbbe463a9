"""Shared by the parity tests: run the same scene through the oracle (CPU restatement of the
reference loops) and through the C ABI (CUDA), and compare with the gates of SURVEY.md §8d:
draw lists / flags bit-exact, float outputs within 2 ulp (0 expected: same operation order, no
contraction on either side)."""
import numpy as np

ULP_TOL = 2  # BASELINE.json north_star: world matrices and bounds within 2 ulp


def ulp_diff(a, b):
    """max distance in units in the last place between two float32 arrays (NaN == NaN, +0 == -0)."""
    a = np.ascontiguousarray(a, np.float32).ravel()
    b = np.ascontiguousarray(b, np.float32).ravel()
    if a.size == 0:
        return 0
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, np.int64(-(2 ** 31)) - ia, ia)
    ib = np.where(ib < 0, np.int64(-(2 ** 31)) - ib, ib)
    d = np.abs(ia - ib)
    both_nan = np.isnan(a) & np.isnan(b)
    d[both_nan] = 0
    return int(d.max())


def run_oracle(orc, scene, views, mode="aabb", omp=False):
    n = scene["num_entities"]
    o = orc.OracleScene(scene["parents"][:n], scene["local_matrices"][:n], scene["entities"][:n],
                        scene["state_flags"][:n], scene["bounding_volumes"][:n],
                        world_matrices=scene.get("world_matrices"), pos_extent=scene.get("pos_extent"),
                        has_rb_matrix=scene.get("has_rb_matrix"), transforms=scene.get("transforms"))
    if views is None or len(views) == 0:
        o.update_scene()
        lists = []
    else:
        lists, _ = o.frame(views, mode=mode, omp=omp)
    return dict(world=o.world, bv=o.bv, pos_extent=o.pos_extent, entities=o.entities, state_flags=o.state_flags,
                extents=o.renderable_extents, lists=lists)


def upload_scene(dev, scene):
    from pmtech_b200 import capi
    n = scene["num_entities"]
    keep = [dev.upload(capi.PARENTS, scene["parents"][:n]),
            dev.upload(capi.LOCAL_MATRICES, scene["local_matrices"][:n]),
            dev.upload(capi.ENTITIES, scene["entities"][:n]),
            dev.upload(capi.STATE_FLAGS, scene["state_flags"][:n]),
            dev.upload(capi.BOUNDING_VOLUMES, scene["bounding_volumes"][:n])]
    if scene.get("has_rb_matrix") is not None:
        keep.append(dev.upload(capi.HAS_RB_MATRIX, scene["has_rb_matrix"][:n]))
    if scene.get("world_matrices") is not None:
        keep.append(dev.upload(capi.WORLD_MATRICES, scene["world_matrices"][:n]))
    if scene.get("pos_extent") is not None:
        keep.append(dev.upload(capi.POS_EXTENT, scene["pos_extent"][:n]))
    if scene.get("transforms") is not None:
        keep.append(dev.upload(capi.TRANSFORMS, scene["transforms"][:n]))
    dev.synchronize()
    return keep


def download_outputs(dev, n, nv):
    from pmtech_b200 import capi
    return dict(world=dev.download(capi.WORLD_MATRICES, 0, n), bv=dev.download(capi.BOUNDING_VOLUMES, 0, n),
                pos_extent=dev.download(capi.POS_EXTENT, 0, n), entities=dev.download(capi.ENTITIES, 0, n),
                state_flags=dev.download(capi.STATE_FLAGS, 0, n), extents=dev.renderable_extents(),
                lists=dev.lists(nv) if nv else [])


def run_device(scene, views, mode="aabb", fused=True, capacity=None, hierarchy=None, tail=None):
    """fused=True: pmb_frame; False: pmb_update_scene then pmb_cull_views (the two-call form).
    hierarchy: None (auto) / "chains" / "wavefront" (PMB_OPT_HIERARCHY); tail: None (default) / False / True
    (PMB_OPT_FUSED_TAIL: the whole frame as one launch for scenes up to 1 Mi entities)."""
    from pmtech_b200 import capi
    from pmtech_b200.scene import DeviceScene
    n = scene["num_entities"]
    nv = 0 if views is None else len(views)
    with DeviceScene(capacity or max(n, 1)) as dev:
        if hierarchy is not None:
            dev.set_option(capi.OPT_HIERARCHY, {"auto": 0, "chains": 1, "wavefront": 2}[hierarchy])
        if tail is not None:
            dev.set_option(capi.OPT_FUSED_TAIL, 2 if tail else 0)
        upload_scene(dev, scene)
        if nv == 0:
            dev.update_scene(n)
        elif fused:
            dev.frame(n, views, mode)
        else:
            dev.update_scene(n)
            dev.cull_views(n, views, mode)
        return download_outputs(dev, n, nv)


def assert_parity(got, want, what=""):
    assert len(got["lists"]) == len(want["lists"]), what
    for v, (g, w) in enumerate(zip(got["lists"], want["lists"])):
        assert g.shape == w.shape, f"{what} view {v}: {g.shape[0]} visible, oracle {w.shape[0]}"
        assert np.array_equal(g, w), f"{what} view {v}: draw list differs"
    assert np.array_equal(got["entities"], want["entities"]), f"{what}: entity flags differ"
    assert np.array_equal(got["state_flags"], want["state_flags"]), f"{what}: state flags differ"
    worst = {}
    for key in ("world", "bv", "pos_extent", "extents"):
        worst[key] = ulp_diff(got[key], want[key])
        assert worst[key] <= ULP_TOL, f"{what}: {key} differs by {worst[key]} ulp (tolerance {ULP_TOL})"
    return worst
