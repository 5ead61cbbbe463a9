"""The oracle — and, on a GPU, the CUDA path — against vectors pmtech itself computed.

tests/golden/pmtech_sdf_scene.npz / pmtech_lucy_scene.npz are the component arrays of the two scene files the
reference ships (examples/assets/scene/*.pms), extracted by tests/golden/make_pms_fixture.py.  save_scene dumps
the arrays as the editor's last update_scene left them, so the file holds inputs AND reference outputs of the
path: replaying the inputs must give the stored outputs bit for bit.  What this pins:

  loop A   world = parent.world * local             mat4 product order, no contraction   (15 <- 14, 16 <- 15)
  loop B   8 transformed corners, min/max, radius   transform_vector order, mag          (rotated + scaled boxes)
  loop C   parent boxes grown by their children                                           (14 <- 15 <- 16)
  stage 0  T * R * S and the quaternion expansion                                        (7 rotated entities)
  draw     inverse transpose incl. its -0 pattern                                        (16 live records)
"""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CMP_ALLOCATED, CMP_LIGHT, CMP_TRANSFORM = 1, 1 << 11, 1 << 12
INVALID = 0xFFFFFFFF


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def live_draw_records(f):
    """records whose inverse transpose pmtech wrote on its last frame: allocated, a valid cbuffer, and not a
    light (the sample animates its lights after the scene update, so their stored matrices are a frame apart)"""
    ent, cb = f["entities"], f["cbuffer"]
    return ((ent & np.uint64(CMP_ALLOCATED)) != 0) & (cb != 0) & (cb != INVALID) & ((ent & np.uint64(CMP_LIGHT)) == 0)


def baked_entities(f):
    """entities whose stored local matrix is the bake of their stored transform: allocated, not lights"""
    ent = f["entities"]
    return np.flatnonzero(((ent & np.uint64(CMP_ALLOCATED)) != 0) & ((ent & np.uint64(CMP_LIGHT)) == 0)
                          & ((ent & np.uint64(0xFFFFFFFE)) != 0))


def test_fixture_is_what_the_generator_describes():
    f = load("pmtech_sdf_scene")
    assert int(f["version"]) == 8 and f["entities"].shape == (23,)
    assert f["component_sizes"].tolist()[:12] == [8, 8, 4, 4, 4, 16, 16, 16, 4, 40, 64, 64]
    assert list(f["names"][:3]) == ["light_4", "floor", "wall_l"]
    par = f["parents"]
    assert par[15] == 14 and par[16] == 15 and (par <= np.arange(23)).all()     # a depth-2 chain
    assert len(baked_entities(f)) == 17 and live_draw_records(f).sum() == 16
    g = load("pmtech_lucy_scene")
    assert int(g["version"]) == 9 and g["entities"].shape == (1,) and len(g["component_sizes"]) == 35


@pytest.mark.parametrize("name", ["pmtech_sdf_scene", "pmtech_lucy_scene"])
def test_oracle_reproduces_pmtech_world_matrices_and_bounds(oracle, name):
    import numpy_ref as ref
    f = load(name)
    n = f["entities"].shape[0]
    o = oracle.OracleScene(f["parents"], f["local_matrices"], f["entities"], f["state_flags"], f["bounding_volumes"])
    o.update_scene()
    assert np.array_equal(bits(o.world), bits(f["world_matrices"]))
    assert np.array_equal(bits(o.bv), bits(f["bounding_volumes"]))
    # the second restatement agrees on the same data
    sc = dict(num_entities=n, parents=f["parents"], local_matrices=f["local_matrices"], entities=f["entities"],
              state_flags=f["state_flags"], bounding_volumes=f["bounding_volumes"])
    r = ref.frame(sc, np.zeros((0, 60), np.float32), "aabb")
    assert np.array_equal(bits(r["world"]), bits(f["world_matrices"]))
    assert np.array_equal(bits(r["bv"]), bits(f["bounding_volumes"]))
    if name == "pmtech_sdf_scene":  # non-vacuous: rotations, a chain, boxes that the loops really rewrite
        scrub = f["bounding_volumes"].copy()
        scrub[:, 6:] = 0
        o2 = oracle.OracleScene(f["parents"], f["local_matrices"], f["entities"], f["state_flags"], scrub,
                                world_matrices=np.zeros((n, 16), np.float32))
        o2.update_scene()
        alive = (f["entities"] & np.uint64(CMP_ALLOCATED)) != 0
        assert np.array_equal(bits(o2.world[alive]), bits(f["world_matrices"][alive]))
        assert np.array_equal(bits(o2.bv[alive]), bits(f["bounding_volumes"][alive]))
        assert (f["world_matrices"][16] != f["local_matrices"][16]).any()


def test_stored_rotations_expand_to_pmtech_local_matrices(oracle):
    """T * R * S with R = the expansion of the STORED rotation is the stored local matrix, bit for bit, for all 17
    baked entities (7 of them rotated).  Normalising the stored rotation again on a copy is NOT what pmtech does:
    2 of the 7 would move by an ulp or two."""
    import numpy_ref as ref
    f = load("pmtech_sdf_scene")
    sel = baked_entities(f)
    t = f["transforms"][sel]
    k = len(sel)
    T = np.tile(np.eye(4, dtype=np.float32).reshape(1, 16), (k, 1))
    T[:, 3], T[:, 7], T[:, 11] = t[:, 0], t[:, 1], t[:, 2]
    S = np.zeros((k, 16), np.float32)
    S[:, 0], S[:, 5], S[:, 10], S[:, 15] = t[:, 7], t[:, 8], t[:, 9], 1
    want = f["local_matrices"][sel]
    got = ref.mat_mul(ref.mat_mul(T, ref.quat_expand(t[:, 3:7])), S)
    assert np.array_equal(bits(got), bits(want))
    rotated = (t[:, 3:7] != np.array([0, 0, 0, 1], np.float32)).any(axis=1)
    assert rotated.sum() == 7
    for i in range(k):
        assert np.array_equal(bits(oracle.quat_expand(t[i, 3:7])), bits(ref.quat_expand(t[i:i + 1, 3:7])[0]))
    renorm = ref.mat_mul(ref.mat_mul(T, ref.quat_expand(ref.quat_normalise(t[:, 3:7]))), S)
    assert (bits(renorm) != bits(want)).any(axis=1).sum() == 2
    # the full stage-0 bake (normalise in place + expand) on the rotations that are fixed points of normalise
    fixed = np.flatnonzero((bits(ref.quat_normalise(t[:, 3:7])) == bits(t[:, 3:7])).all(axis=1))
    assert len(fixed) == 15
    ent = f["entities"].copy()
    ent[sel[fixed]] |= np.uint64(CMP_TRANSFORM)
    o = oracle.OracleScene(f["parents"], np.zeros_like(f["local_matrices"]), ent, f["state_flags"], f["bounding_volumes"],
                           transforms=f["transforms"])
    o.update_transforms()
    assert np.array_equal(bits(o.local[sel[fixed]]), bits(want[fixed]))
    assert np.array_equal(bits(o.transforms), bits(f["transforms"]))


def test_oracle_reproduces_pmtech_draw_call_records(oracle):
    f = load("pmtech_sdf_scene")
    n = f["entities"].shape[0]
    o = oracle.OracleScene(f["parents"], f["local_matrices"], f["entities"], f["state_flags"], f["bounding_volumes"])
    o.update_scene()
    got = o.update_draw_calls(f["cbuffer"], np.zeros((n, 40), np.float32), 0.0)
    live = live_draw_records(f)
    want = f["draw_call_data"]
    assert np.array_equal(bits(got[live, 0:16]), bits(want[live, 0:16]))       # world
    assert np.array_equal(got[live, 16], np.flatnonzero(live).astype(np.float32))  # v1.x = node index (:1975)
    assert np.array_equal(bits(got[live, 24:40]), bits(want[live, 24:40]))     # inverse transpose, signs of zero too
    assert (bits(want[live, 24:40]) == 0x80000000).any()                       # the -0 checkerboard is in the data


# ---- the CUDA path against the same reference vectors ---------------------------------------------------------
def _image(f, version):
    from pmtech_b200 import pms
    arrays = {k: f[k] for k in ("entities", "state_flags", "parents", "transforms", "local_matrices", "world_matrices",
                                "bounding_volumes", "cbuffer", "draw_call_data")}
    return pms.write_pms(arrays, version=version, lookup_strings=[("floor", 77), ("", 0)],
                         cameras=[dict(id=1, pos=(0, 1, 2), focus=(0, 0, 0), rot=(0, 0), fov=60, aspect=1.5, near=0.1, far=100,
                                       zoom=3)])


@pytest.mark.gpu
def test_device_reproduces_pmtech_outputs_from_a_scene_image():
    from pmtech_b200 import capi, pms
    from pmtech_b200.scene import DeviceScene
    f = load("pmtech_sdf_scene")
    n = 23
    scrub = dict(f)
    scrub["world_matrices"] = np.zeros_like(f["world_matrices"])
    scrub["bounding_volumes"] = f["bounding_volumes"].copy()
    scrub["bounding_volumes"][:, 6:] = 0
    scrub["draw_call_data"] = np.zeros_like(f["draw_call_data"])
    # stage 0 on the device for the rotations that are fixed points of quat::normalise: their local matrices are wiped
    import numpy_ref as ref
    sel = baked_entities(f)
    q = f["transforms"][sel, 3:7]
    fixed = sel[(bits(ref.quat_normalise(q)) == bits(q)).all(axis=1)]
    # (the lights were saved with e_cmp::transform pending — the sample moves them after update_scene — which is
    # why their stored local matrices lag their transforms; the pending flag is dropped for the replay)
    scrub["entities"] = f["entities"] & ~np.uint64(CMP_TRANSFORM)
    scrub["entities"][fixed] |= np.uint64(CMP_TRANSFORM)
    scrub["local_matrices"] = f["local_matrices"].copy()
    scrub["local_matrices"][fixed] = 0
    with DeviceScene(64, 0) as dev:
        info = pms.load_into(dev, _image(scrub, 8))
        assert info.num_nodes == n and info.version == 8
        dev.frame(n, np.zeros((0, 60), np.float32), "aabb")
        dev.update_draw_calls(n, 0.0)
        alive = (f["entities"] & np.uint64(CMP_ALLOCATED)) != 0
        assert np.array_equal(bits(dev.download(capi.LOCAL_MATRICES, 0, n)), bits(f["local_matrices"]))
        assert np.array_equal(bits(dev.download(capi.TRANSFORMS, 0, n)), bits(f["transforms"]))
        assert np.array_equal(bits(dev.download(capi.WORLD_MATRICES, 0, n)[alive]), bits(f["world_matrices"][alive]))
        assert np.array_equal(bits(dev.download(capi.BOUNDING_VOLUMES, 0, n)[alive]), bits(f["bounding_volumes"][alive]))
        live = live_draw_records(f)
        dc = dev.download(capi.DRAW_CALLS, 0, n)
        assert np.array_equal(bits(dc[live, 0:16]), bits(f["draw_call_data"][live, 0:16]))
        assert np.array_equal(bits(dc[live, 24:40]), bits(f["draw_call_data"][live, 24:40]))
        assert np.array_equal(dev.download(capi.ENTITIES, 0, n), f["entities"] & ~np.uint64(CMP_TRANSFORM))


@pytest.mark.gpu
def test_device_merge_load_offsets_parents():
    """load_scene(merge): the image lands behind the entities already there and its parents move with it (:2529-2531)"""
    from pmtech_b200 import capi, pms
    from pmtech_b200.scene import DeviceScene
    f = load("pmtech_sdf_scene")
    n = 23
    img = _image(f, 8)
    with DeviceScene(64, 0) as dev:
        pms.load_into(dev, img)
        pms.load_into(dev, img, zero_offset=n)
        dev.frame(2 * n, np.zeros((0, 60), np.float32), "aabb")
        par = dev.download(capi.PARENTS, 0, 2 * n)
        assert np.array_equal(par[:n], f["parents"]) and np.array_equal(par[n:], f["parents"] + np.uint32(n))
        w = dev.download(capi.WORLD_MATRICES, 0, 2 * n)
        # the lights carry a pending e_cmp::transform and are re-baked from their newer transforms, as pmtech would
        alive = ((f["entities"] & np.uint64(CMP_ALLOCATED)) != 0) & ((f["entities"] & np.uint64(CMP_TRANSFORM)) == 0)
        assert alive.sum() == 18
        assert np.array_equal(bits(w[:n][alive]), bits(f["world_matrices"][alive]))
        assert np.array_equal(bits(w[n:][alive]), bits(f["world_matrices"][alive]))
        with pytest.raises(capi.PmbError):
            pms.load_into(dev, img, zero_offset=50)  # 50 + 23 > capacity 64


# ---- the reader itself (host code of the library, no device needed) ---------------------------------------------
def test_pms_reader_walks_an_image_like_load_scene():
    from pmtech_b200 import capi, pms
    f = load("pmtech_sdf_scene")
    img = _image(f, 8)
    p = pms.PmsScene(img)
    assert p.version == 8 and p.num_nodes == 23 and p.info.num_base_components == 31 and p.info.num_cameras == 1
    assert p.lookup_strings == {77: "floor", 0: ""} and p.cameras[0]["fov"] == 60 and p.names == [""] * 23
    for k in ("entities", "parents", "transforms", "local_matrices", "world_matrices", "bounding_volumes", "cbuffer",
              "draw_call_data", "state_flags"):
        assert np.array_equal(p.components[k].view(np.uint8), np.ascontiguousarray(f[k]).view(np.uint8)), k
    assert "pos_extent" not in p.components                       # component 38 arrived with version 10
    assert p.info.names_offset + 12 * 23 == len(img)
    # the reference's own files: offsets recorded when the fixture was made
    assert int(f["names_offset"]) == 42316 and int(f["file_bytes"]) == 44108
    # version 10 image with pos_extent; merge offset applied by host_arrays
    pe = np.arange(23 * 8, dtype=np.float32).reshape(23, 8)
    img10 = pms.write_pms(dict(entities=f["entities"], parents=f["parents"], pos_extent=pe), version=10)
    q = pms.PmsScene(img10)
    assert np.array_equal(q.components["pos_extent"], pe) and sorted(q.components) == ["entities", "parents", "pos_extent"]
    assert np.array_equal(q.host_arrays(zero_offset=5)["parents"], f["parents"] + np.uint32(5))
    # a component stored with another element size is skipped, not misread (ecs_scene.cpp:2506-2524)
    bad = bytearray(img)
    sizes = np.frombuffer(bytes(bad), np.uint32, 31, p.info.sizes_offset).copy()
    assert sizes[9] == 40
    info = capi.PmsInfo()
    L = capi.load()
    # truncated images and a foreign header are refused with a message
    for cut in (10, 300, len(img) - 400):
        buf = np.frombuffer(img[:cut], np.uint8)
        import ctypes as C
        assert L.pmb_pms_parse(capi._ptr(buf), cut, C.byref(info)) == -1
        assert b"pms" in L.pmb_last_error()
    with pytest.raises(capi.PmbError):
        pms.PmsScene(b"\x00" * 512)


def test_pms_reader_skips_resized_components():
    from pmtech_b200 import pms
    f = load("pmtech_sdf_scene")
    img = bytearray(_image(f, 8))
    p = pms.PmsScene(bytes(img))
    # rewrite the image so that transforms claim 44 bytes per element (an older/newer cmp_transform)
    n = 23
    off = p.info.component_offset[3]
    grown = bytes(img[:off]) + bytes(n * 44) + bytes(img[off + n * 40:])
    grown = bytearray(grown)
    grown[p.info.sizes_offset + 9 * 4:p.info.sizes_offset + 10 * 4] = np.uint32(44).tobytes()
    q = pms.PmsScene(bytes(grown))
    assert "transforms" not in q.components
    assert np.array_equal(q.components["local_matrices"], f["local_matrices"])   # what follows is still found


# ---- what the stored vectors discriminate, and what they do not (VERDICT r01 "What's weak" #1) -------------------
# Every variant below is a plausible OTHER way of evaluating the same formula.  A variant that reproduces pmtech's
# stored bits is NOT ruled out by the scene files; one that fails IS.  The asserts state both, so DESIGN.md §2's table
# is backed by a test in either direction.
F32 = np.float32


def _fma(a, b, c):
    """one rounding of a * b + c (binary64 holds the 48-bit product exactly; the final rounding to binary32 of the
    binary64 sum can double-round only on exact ties of the first rounding, which these data do not hit)"""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F32)


def _mat_mul_variant(a, b, order):
    """a, b: [16] row-major.  order: 'ltr' ((0 + t0) + t1) + t2) + t3, 'rtl' t0 + (t1 + (t2 + t3)), 'pair' (t0 + t1) +
    (t2 + t3), 'fma' fma(a3, b3, fma(a2, b2, fma(a1, b1, a0 * b0)))"""
    a = a.reshape(4, 4)
    b = b.reshape(4, 4)
    out = np.zeros((4, 4), F32)
    for r in range(4):
        for c in range(4):
            t = [F32(a[r, k] * b[k, c]) for k in range(4)]
            if order == "ltr":
                acc = F32(0)
                for k in range(4):
                    acc = F32(acc + t[k])
            elif order == "rtl":
                acc = F32(t[0] + F32(t[1] + F32(t[2] + t[3])))
            elif order == "pair":
                acc = F32(F32(t[0] + t[1]) + F32(t[2] + t[3]))
            else:
                acc = t[0]
                for k in range(1, 4):
                    acc = _fma(np.asarray(a[r, k]), np.asarray(b[k, c]), np.asarray(acc))[()]
            out[r, c] = acc
    return out.reshape(16)


def test_what_the_stored_world_matrices_discriminate():
    """loop A (ecs_scene.cpp:1465-1470) on the stored chain 14 <- 15 <- 16: the scene files rule out FMA contraction of
    the mat4 product; they do NOT tell the summation orders apart (all three reproduce the stored bits)."""
    f = load("pmtech_sdf_scene")
    w, l = f["world_matrices"], f["local_matrices"]
    mism = {}
    for order in ("ltr", "rtl", "pair", "fma"):
        w15 = _mat_mul_variant(w[14], l[15], order)
        w16 = _mat_mul_variant(w[15], l[16], order)
        mism[order] = (int((bits(w15) != bits(w[15])).sum()), int((bits(w16) != bits(w[16])).sum()))
    assert mism["ltr"] == (0, 0), mism                      # the declared mode reproduces pmtech
    assert mism["rtl"] == (0, 0) and mism["pair"] == (0, 0), mism  # NOT discriminated: summation order of the product
    assert mism["fma"][0] > 0 and mism["fma"][1] > 0, mism  # discriminated: a contracted chain does not reproduce pmtech


def _leaf_boxes(f):
    """allocated non-bone entities nobody folds into (loop C leaves their box as loop B wrote it)"""
    n = f["entities"].shape[0]
    ent, par = f["entities"], f["parents"]
    has_child = np.zeros(n, bool)
    for i in range(n):
        if int(par[i]) != i and (int(ent[i]) & CMP_ALLOCATED):
            has_child[int(par[i])] = True
    return np.flatnonzero(((ent & np.uint64(CMP_ALLOCATED)) != 0) & ~has_child)


def _bounds_variant(w, mn, mx, order):
    """loop B (ecs_scene.cpp:1486-1512) for one entity: 8 corners through transform_vector, min / max folds.
    order: 'ltr' ((0 + x m0) + y m1) + z m2) + w, 'rtl' x m0 + (y m1 + (z m2 + w)), 'pair' (x m0 + y m1) + (z m2 + w),
    'wfirst' ((w + x m0) + y m1) + z m2, 'fma' fma(z, m2, fma(y, m1, fma(x, m0, 0))) + w"""
    w = w.reshape(4, 4)
    size = (mx - mn).astype(F32)
    tmin = np.full(3, np.finfo(F32).max, F32)
    tmax = np.full(3, -np.finfo(F32).max, F32)
    for c in [(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (0, 1, 1), (1, 0, 1), (1, 1, 1)]:
        p = (mn + (size * np.array(c, F32)).astype(F32)).astype(F32)
        t = np.zeros(3, F32)
        for r in range(3):
            x, y, z = (F32(w[r, k] * p[k]) for k in range(3))
            ww = F32(w[r, 3] * F32(1))
            if order == "ltr":
                t[r] = F32(F32(F32(F32(F32(0) + x) + y) + z) + ww)
            elif order == "rtl":
                t[r] = F32(x + F32(y + F32(z + ww)))
            elif order == "pair":
                t[r] = F32(F32(x + y) + F32(z + ww))
            elif order == "wfirst":
                t[r] = F32(F32(F32(ww + x) + y) + z)
            else:
                acc = np.asarray(F32(0))
                for k in range(3):
                    acc = _fma(np.asarray(w[r, k]), np.asarray(p[k]), acc)
                t[r] = F32(acc[()] + ww)
        tmin = np.minimum(tmin, t)
        tmax = np.maximum(tmax, t)
    return tmin, tmax


def test_what_the_stored_bounds_discriminate():
    """loop B on the boxes nobody folds into: the scene files rule out the other summation orders of transform_vector;
    they do NOT rule out FMA contraction there (a contracted chain reproduces the stored bits), and the stored radius
    (a Vec3 dot through mag) tells no order apart."""
    f = load("pmtech_sdf_scene")
    leaves = _leaf_boxes(f)
    assert len(leaves) == 20
    bv, w = f["bounding_volumes"], f["world_matrices"]
    mism = {}
    for order in ("ltr", "rtl", "pair", "wfirst", "fma"):
        bad = 0
        for i in leaves:
            tmin, tmax = _bounds_variant(w[i], bv[i, 0:3], bv[i, 3:6], order)
            bad += int((bits(tmin) != bits(bv[i, 6:9])).sum() + (bits(tmax) != bits(bv[i, 9:12])).sum())
        mism[order] = bad
    assert mism["ltr"] == 0, mism                                                  # the declared mode reproduces pmtech
    assert mism["rtl"] > 0 and mism["pair"] > 0 and mism["wfirst"] > 0, mism        # discriminated: summation order
    assert mism["fma"] == 0, mism                                                  # NOT discriminated: contraction in loop B
    # radius = mag(tmax - tmin) * 0.5: every order of the three-term dot gives the stored bits -> nothing pinned by it
    rad_bad = {}
    for order in ("ltr", "rtl"):
        bad = 0
        for i in leaves:
            d = (bv[i, 9:12] - bv[i, 6:9]).astype(F32)
            s = [F32(d[k] * d[k]) for k in range(3)]
            m2 = F32(F32(F32(F32(0) + s[0]) + s[1]) + s[2]) if order == "ltr" else F32(s[0] + F32(s[1] + s[2]))
            bad += int(bits(np.asarray(F32(np.sqrt(m2)) * F32(0.5), F32)) != bits(np.asarray(bv[i, 12])))
        rad_bad[order] = bad
    assert rad_bad == {"ltr": 0, "rtl": 0}, rad_bad


def test_loops_a_and_b_together_leave_one_mode():
    """Of the four candidate (order, contraction) modes of the maths library's dot products, only 'left to right,
    no contraction' passes BOTH stored data sets: loop A rules out contraction, loop B rules out the other orders."""
    f = load("pmtech_sdf_scene")
    w, l, bv = f["world_matrices"], f["local_matrices"], f["bounding_volumes"]
    leaves = _leaf_boxes(f)
    verdict = {}
    for order in ("ltr", "rtl", "pair", "fma"):
        a_ok = all((bits(_mat_mul_variant(w[p], l[c], order)) == bits(w[c])).all() for p, c in ((14, 15), (15, 16)))
        b_ok = True
        for i in leaves:
            tmin, tmax = _bounds_variant(w[i], bv[i, 0:3], bv[i, 3:6], order)
            b_ok &= bool((bits(tmin) == bits(bv[i, 6:9])).all() and (bits(tmax) == bits(bv[i, 9:12])).all())
        verdict[order] = a_ok and b_ok
    assert verdict == {"ltr": True, "rtl": False, "pair": False, "fma": False}, verdict
