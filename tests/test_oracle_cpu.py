"""CPU suite, part 1: the oracle itself.

The reference holds no golden vectors for this path (its tests are golden images
through a graphics API, and the arithmetic lives in the absent third_party/maths submodule), so
the oracle is pinned by what can be checked without it:
  * closed-form cases where every product is exact, so any correct operation order gives the
    same bits (translations, power-of-two scales, 90-degree rotations);
  * a second restatement written independently in numpy (tests/numpy_ref.py), bit for bit;
  * a float64 shadow computation bounding the float32 error;
  * the invariants of the reference loops (lists ascending, culled within filtered, boxes ordered);
  * the committed fixtures under tests/golden/ (made by tests/golden/make_golden.py).
"""
import os

import numpy as np
import pytest

import numpy_ref
from helpers import run_oracle, ulp_diff

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def cams(oracle):
    return oracle.OracleCamera


def _trs(t, s=(1, 1, 1)):
    m = np.eye(4, dtype=np.float32)
    m[0, 0], m[1, 1], m[2, 2] = s
    m[:3, 3] = t
    return m.reshape(16)


def test_mat4_mul_exact_cases(oracle):
    a = _trs((1, 2, 3), (2, 4, 8))
    b = _trs((-4, 0.5, 16), (0.5, 0.25, 2))
    got = oracle.mat4_mul(a, b).reshape(4, 4)
    want = (a.reshape(4, 4).astype(np.float64) @ b.reshape(4, 4).astype(np.float64)).astype(np.float32)
    assert np.array_equal(got, want)
    # column-vector convention: translation lives in m[3], m[7], m[11] (physics_bullet.cpp:310-315)
    assert got[0, 3] == 1 + 2 * -4 and got[1, 3] == 2 + 4 * 0.5 and got[2, 3] == 3 + 8 * 16
    rot90 = np.array([[0, -1, 0, 0], [1, 0, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], np.float32).reshape(16)
    p = oracle.transform_vector(oracle.mat4_mul(_trs((10, 20, 30)), rot90), np.array([1, 2, 3], np.float32))
    assert p.tolist() == [8.0, 21.0, 33.0]


def test_mat4_mul_order_is_left_to_right(oracle):
    """values chosen so that (a0*b0 + a1*b1) + a2*b2 + a3*b3 differs from any other association"""
    a = np.zeros(16, np.float32)
    b = np.zeros(16, np.float32)
    a[0:4] = [1.0, 1.0, 1.0, 1.0]
    b[0], b[4], b[8], b[12] = 2.0 ** 24, 1.0, -(2.0 ** 24), 1.0
    got = oracle.mat4_mul(a, b)[0]
    f = np.float32
    want = f(f(f(f(0) + f(2.0 ** 24)) + f(1.0)) + f(-(2.0 ** 24))) + f(1.0)
    assert got == want == 1.0  # ((2^24 + 1) - 2^24) + 1 with the +1 lost in the first add


def test_plane_distance_and_cull_sign_convention(oracle, cams):
    cam = cams().create_perspective(60.0, 16.0 / 9.0, 0.1, 100.0)
    cam.look_at((0, 0, 0), (0, 0, -1)).update_frustum()
    fr = cam.frustum
    n, p = fr[0:18].reshape(6, 3), fr[18:36].reshape(6, 3)
    inside = np.array([0, 0, -10], np.float32)
    for k in range(6):
        # outward normals: a point inside the frustum is on the negative side of all six planes
        assert float(np.dot(inside - p[k], n[k])) < 0
    # maths::plane_distance is such that dot(x, n) + pd is the signed distance (ecs_cull.cpp:37-40)
    for k in range(6):
        pd = C_plane_distance(oracle, p[k], n[k])
        assert abs((np.dot(inside, n[k]) + pd) - np.dot(inside - p[k], n[k])) < 1e-3


def C_plane_distance(oracle, x0, xn):
    x0 = np.ascontiguousarray(x0, np.float32)
    xn = np.ascontiguousarray(xn, np.float32)
    return float(oracle.lib().orc_plane_distance(x0.ctypes.data, xn.ctypes.data))


def test_closed_form_scene(oracle, cams):
    """pure translations and power-of-two scales: every product and sum is exact"""
    n = 64
    parents = np.arange(n, dtype=np.uint32)
    parents[1:] = (np.arange(1, n) - 1) // 2  # binary tree
    local = np.stack([_trs((i % 5 - 2, (i * 3) % 7 - 3, i % 3), (2.0 if i % 2 else 0.5,) * 3) for i in range(n)])
    ent = np.full(n, 0x13, np.uint64)
    sf = np.zeros(n, np.uint64)
    bv = np.zeros((n, 13), np.float32)
    bv[:, 0:3], bv[:, 3:6] = -1, 1
    sc = dict(num_entities=n, parents=parents, local_matrices=local, entities=ent, state_flags=sf, bounding_volumes=bv)
    got = run_oracle(oracle, sc, None)
    # float64 chain
    w = np.zeros((n, 4, 4))
    for i in range(n):
        l = local[i].reshape(4, 4).astype(np.float64)
        w[i] = l if parents[i] == i else w[parents[i]] @ l
    assert np.array_equal(got["world"], w.reshape(n, 16).astype(np.float32))
    for i in range(n):
        c = w[i][:3, 3]
        s = w[i][0, 0]
        assert np.array_equal(got["pos_extent"][i, 0:3], c.astype(np.float32))
        assert np.array_equal(got["pos_extent"][i, 4:7], np.full(3, s, np.float32))


@pytest.mark.parametrize("n,seed,mode", [(1, 1, "aabb"), (257, 2, "aabb"), (3000, 3, "aabb"), (3000, 4, "sphere")])
def test_oracle_matches_independent_numpy_restatement(oracle, cams, n, seed, mode):
    from pmtech_b200 import synth
    sc = synth.mixed_scene(n=n, seed=seed)
    views = synth.mixed_views(cams)
    want = numpy_ref.frame(sc, views, mode)
    got = run_oracle(oracle, sc, views, mode=mode)
    for k in ("world", "bv", "pos_extent", "extents"):
        assert ulp_diff(got[k], want[k]) == 0, k
        assert np.array_equal(np.asarray(got[k]).view(np.uint32), np.asarray(want[k]).view(np.uint32)), k
    assert np.array_equal(got["entities"], want["entities"])
    assert np.array_equal(got["state_flags"], want["state_flags"])
    for v in range(len(views)):
        assert np.array_equal(got["lists"][v], want["lists"][v]), v


def test_oracle_omp_variant_is_bit_identical(oracle, cams):
    from pmtech_b200 import synth
    sc = synth.mixed_scene(n=20000, seed=8)
    views = synth.mixed_views(cams)
    a = run_oracle(oracle, sc, views, omp=False)
    b = run_oracle(oracle, sc, views, omp=True)
    for k in ("world", "bv", "pos_extent", "extents"):
        assert np.array_equal(np.asarray(a[k]).view(np.uint32), np.asarray(b[k]).view(np.uint32)), k
    for v in range(len(views)):
        assert np.array_equal(a["lists"][v], b["lists"][v])


def test_float64_shadow_bounds_the_error(oracle):
    from pmtech_b200 import synth
    sc = synth.deep_scene(n=1 << 14, max_depth=16, seed=2, block=4096)
    got = run_oracle(oracle, sc, None)
    n = sc["num_entities"]
    w = np.zeros((n, 4, 4))
    par = sc["parents"]
    loc = sc["local_matrices"].reshape(n, 4, 4).astype(np.float64)
    for i in range(n):
        w[i] = loc[i] if par[i] == i else w[par[i]] @ loc[i]
    err = np.abs(got["world"].reshape(n, 4, 4) - w)
    scale = np.maximum(np.abs(w).max(axis=(1, 2), keepdims=True), 1.0)
    assert (err / scale).max() < 16 * 2 * np.finfo(np.float32).eps * 4  # depth x 4-term dots


def test_invariants(oracle, cams):
    from pmtech_b200 import synth
    sc = synth.city_scene(100_000, seed=12)
    views = synth.city_views(cams)
    got = run_oracle(oracle, sc, views)
    assert (sc["parents"] <= np.arange(sc["num_entities"])).all()  # ecs_utilities.cpp:408-415
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"])
    o.update_scene()
    filt = o.filter_entities()
    assert (np.diff(filt.astype(np.int64)) > 0).all()
    for l in got["lists"]:
        assert (np.diff(l.astype(np.int64)) > 0).all()
        assert np.isin(l, filt).all()
    bv, pe = got["bv"], got["pos_extent"]
    assert (bv[:, 6:9] <= bv[:, 9:12]).all()
    assert (pe[:, 4:7] >= 0).all()
    # sphere cull keeps at least what the box cull keeps? no: different tests; but both within filter
    for v, fr in enumerate(views[:2]):
        s = o.frustum_cull(fr, filt, "sphere")
        assert np.isin(s, filt).all()
    # culling an explicit sub-list keeps the caller's order (ecs_cull.cpp:29-49)
    sub = filt[::7][::-1].copy()
    out = o.frustum_cull(views[4], sub, "aabb")
    pos = {int(e): i for i, e in enumerate(sub)}
    assert all(pos[int(a)] < pos[int(b)] for a, b in zip(out[:-1], out[1:]))


def test_stretchy_buffer_header(oracle):
    """core/pen/include/data_struct.h:22-72: {int capacity, int count} 8 bytes before the data,
    doubling growth"""
    import ctypes as C
    L = oracle.lib()
    sb = C.c_void_p(None)
    for i in range(100):
        L.orc_sb_push(C.byref(sb), i * 3)
    assert L.orc_sb_count(sb) == 100
    hdr = np.ctypeslib.as_array(C.cast(sb.value - 8, C.POINTER(C.c_int)), shape=(2,))
    assert hdr[1] == 100 and hdr[0] >= 100
    data = np.ctypeslib.as_array(C.cast(sb, C.POINTER(C.c_uint32)), shape=(100,))
    assert data.tolist() == [i * 3 for i in range(100)]
    L.orc_sb_free(sb)
    assert L.orc_sb_count(None) == 0


@pytest.mark.parametrize("name", ["mixed_3000", "torus_views", "rigs_40"])
def test_golden_fixtures(oracle, cams, name):
    """committed fixtures (tests/golden/make_golden.py): the oracle must keep producing them"""
    import make_golden_lib as mg
    path = os.path.join(GOLDEN, name + ".npz")
    assert os.path.exists(path), "run tests/golden/make_golden.py"
    z = np.load(path)
    sc, views, mode = mg.scene_for(name, cams)
    got = run_oracle(oracle, sc, views, mode=mode)
    for k in ("world", "bv", "pos_extent", "extents"):
        assert np.array_equal(np.asarray(got[k]).view(np.uint32), z[k].view(np.uint32)), k
    assert np.array_equal(views, z["views"])
    for v in range(len(views)):
        assert np.array_equal(got["lists"][v], z[f"list{v}"])


# ---- stage 0 (TRS bake), root motion, skin palettes: oracle vs the independent numpy restatement ----
def _rig_with_flags(seed=5, characters=12):
    from pmtech_b200 import synth
    import numpy_ref as ref
    sc = synth.rig_scene(characters, seed=seed)
    rng = np.random.default_rng(seed)
    n = sc["num_entities"]
    flagged = rng.random(n) < 0.7
    sc["entities"] = sc["entities"] | np.where(flagged, np.uint64(ref.CMP_TRANSFORM), np.uint64(0))
    # a few entities are forced to sync with physics: their transform flag is dropped unbaked (:1408-1412)
    sync = rng.random(n) < 0.05
    sc["state_flags"] = sc["state_flags"] | np.where(sync, np.uint64(ref.STATE_SYNC_PHYSICS_TRANSFORM), np.uint64(0))
    # entities that are not re-baked keep whatever local matrix they had: make that visible
    sc["local_matrices"] = sc["local_matrices"].copy()
    sc["local_matrices"][~flagged, 3] += np.float32(0.5)
    return sc


def test_stage0_bake_matches_numpy_and_the_host_trs(oracle):
    import numpy_ref as ref
    sc = _rig_with_flags()
    n = sc["num_entities"]
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           transforms=sc["transforms"])
    o.update_transforms()
    want, t_want = ref.bake_transforms(sc)
    assert np.array_equal(o.local.view(np.uint32), want.view(np.uint32))
    # quat::get_matrix normalises the stored rotation in place: the transforms are an output of stage 0
    assert np.array_equal(o.transforms.view(np.uint32), t_want.view(np.uint32))
    assert (o.transforms.view(np.uint32) != sc["transforms"].view(np.uint32)).any()
    assert not (o.entities & np.uint64(ref.CMP_TRANSFORM)).any()
    assert not (o.state_flags & np.uint64(ref.STATE_SYNC_PHYSICS_TRANSFORM)).any()
    baked = ((sc["entities"] & np.uint64(ref.CMP_TRANSFORM)) != 0) & \
            ((sc["state_flags"] & np.uint64(ref.STATE_SYNC_PHYSICS_TRANSFORM)) == 0)
    assert baked.any() and (~baked).any()
    assert np.array_equal(o.local[~baked], sc["local_matrices"][~baked])
    # T * R * S agrees with the closed-form host bake of the generator to rounding
    from pmtech_b200 import synth
    t = sc["transforms"]
    close = synth.trs_matrices(t[:, 0:3], t[:, 3:7], t[:, 7:10])
    assert np.allclose(o.local[baked], close[baked], rtol=0, atol=2e-6)
    # exact cases: identity rotation, power-of-two scale
    tr = np.array([[1, 2, 3, 0, 0, 0, 1, 2, 4, 0.5]], np.float32)
    o1 = oracle.OracleScene(np.zeros(1, np.uint32), np.zeros((1, 16), np.float32), np.array([1 | (1 << 12)], np.uint64),
                            np.zeros(1, np.uint64), np.zeros((1, 13), np.float32), transforms=tr)
    o1.update_transforms()
    assert o1.local[0].tolist() == [2, 0, 0, 1, 0, 4, 0, 2, 0, 0, 0.5, 3, 0, 0, 0, 1]


def test_root_motion_and_skin_palettes_match_numpy(oracle):
    import numpy_ref as ref
    from pmtech_b200 import synth
    sc = synth.rig_scene(16, seed=9)
    rm = sc["root_motion"]
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           transforms=sc["transforms"])
    o.apply_root_motion(rm["targets"], rm["rotations"], rm["deltas"])
    t_want, e_want = ref.apply_root_motion(sc, rm["targets"], rm["rotations"], rm["deltas"])
    assert np.array_equal(o.transforms.view(np.uint32), t_want.view(np.uint32))
    assert np.array_equal(o.entities, e_want)
    assert ((o.entities[rm["targets"]] & np.uint64(ref.CMP_TRANSFORM)) != 0).all()
    moved = np.abs(o.transforms[rm["targets"], 0:3] - sc["transforms"][rm["targets"], 0:3]).max()
    assert 0 < moved < 0.2
    o.update_scene()
    pal = o.skin_palettes(sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])
    want = ref.skin_palettes(o.world, sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])
    assert np.array_equal(pal.view(np.uint32), want.view(np.uint32))
    assert np.abs(pal).max() > 1


def test_draw_call_packing_matches_numpy_and_inverts(oracle):
    import numpy_ref as ref
    from pmtech_b200 import synth
    sc = synth.mixed_scene(n=3000, seed=12)
    n = sc["num_entities"]
    rng = np.random.default_rng(3)
    ent = sc["entities"].copy()
    ent[rng.random(n) < 0.05] |= np.uint64(1 << 6)    # skinned
    ent[rng.random(n) < 0.03] |= np.uint64(1 << 16)   # pre_skinned
    cb = rng.integers(1, 5000, n).astype(np.uint32)
    cb[rng.random(n) < 0.1] = 0
    cb[rng.random(n) < 0.1] = 0xFFFFFFFF
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], ent, sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc.get("world_matrices"), pos_extent=sc.get("pos_extent"),
                           has_rb_matrix=sc.get("has_rb_matrix"))
    o.update_scene()
    dc0 = rng.standard_normal((n, 40)).astype(np.float32)  # user data in v1.zw / v2 and stale matrices must survive
    got = o.update_draw_calls(cb, dc0.copy(), 1234.5)
    want = ref.update_draw_calls(o.world, o.entities, cb, dc0, 1234.5)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(got[:, 18:24], dc0[:, 18:24])
    live = (cb != 0) & (cb != 0xFFFFFFFF) & ((o.entities & np.uint64(1 << 14)) == 0)
    assert np.array_equal(got[~live, 24:40], dc0[~live, 24:40]) and live.any() and (~live).any()
    # it is the inverse of the transpose: (W^T)^-1 * W^T = I to rounding, on well-conditioned matrices
    w = o.world[live].reshape(-1, 4, 4).astype(np.float64)
    inv = got[live, 24:40].reshape(-1, 4, 4).astype(np.float64)
    ok = np.isfinite(inv).all(axis=(1, 2)) & (np.abs(np.linalg.det(w)) > 1e-3) & (np.abs(w).max(axis=(1, 2)) < 1e4)
    err = np.abs(inv[ok] @ w[ok].transpose(0, 2, 1) - np.eye(4)).max()
    assert ok.sum() > 100 and err < 1e-3, err
    # exact case: a power-of-two scale + translation
    m = np.array([[2, 0, 0, 0, 0, 4, 0, 0, 0, 0, 0.5, 0, 1, 2, 3, 1]], np.float32)
    assert ref.inverse4x4(m)[0].tolist() == [0.5, 0, 0, 0, 0, 0.25, 0, 0, 0, 0, 2, 0, -0.5, -0.5, -6, 1]


def test_golden_stage0_fixture(oracle, cams):
    """tests/golden/rigs_stage0.npz: root motion + bake + frame + palettes + draw calls stay what they were"""
    import make_golden_lib as mg
    z = np.load(os.path.join(GOLDEN, "rigs_stage0.npz"))
    case = mg.stage0_case(cams)
    got = mg.run_stage0_oracle(oracle, case)
    for k in mg.STAGE0_KEYS:
        assert np.array_equal(np.ascontiguousarray(got[k]).view(np.uint32 if got[k].dtype == np.float32 else got[k].dtype),
                              z[k].view(np.uint32 if z[k].dtype == np.float32 else z[k].dtype)), k
    for v in range(len(case[1])):
        assert np.array_equal(got["lists"][v], z[f"list{v}"])


def test_spot_shadow_camera_host_producer_matches_the_oracle(oracle):
    """ecs_scene.cpp:556-570 — the library's host producer (shared source with the device fit) against the oracle's"""
    from pmtech_b200 import ecs
    rng = np.random.default_rng(17)
    for _ in range(20):
        w = np.eye(4, dtype=np.float32)
        w[:3, :3] = np.linalg.qr(rng.standard_normal((3, 3)))[0].astype(np.float32) * rng.uniform(0.5, 3.0, 3).astype(np.float32)
        w[:3, 3] = rng.uniform(-50, 50, 3).astype(np.float32)
        a = oracle.OracleCamera().spot_shadow(w.reshape(16))
        b = ecs.Camera().spot_shadow(w.reshape(16))
        for x, y in ((a.view, b.view), (a.proj, b.proj), (a.frustum, b.frustum)):
            assert np.array_equal(np.asarray(x, np.float32).view(np.uint32), np.asarray(y, np.float32).view(np.uint32))
        assert np.isfinite(a.frustum).all()
