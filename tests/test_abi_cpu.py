"""CPU suite, part 2: the boundary.  The C-ABI library loads without a GPU, exports every symbol
include/pmtech_b200.h declares, refuses to run without a device (no CPU fallback), and its
host-side view producers (camera.cpp restatements) agree bit for bit with the oracle's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "pmtech_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pmb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(lib):
    from pmtech_b200 import capi
    hs = header_symbols()
    assert len(hs) >= 39
    assert sorted(capi.SYMBOLS) == hs, set(hs) ^ set(capi.SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for s in header_symbols():
        assert hasattr(lib, s), f"{s} is declared in include/pmtech_b200.h but not exported"
    assert b"sm_100a" in lib.pmb_version()


def test_struct_layouts_match_the_reference_records(lib):
    from pmtech_b200 import capi
    assert C.sizeof(capi.CameraStruct) == 4 * (3 + 16 + 16 + 60)  # pmb_camera
    assert capi.FRUSTUM_FLOATS * 4 == 240                        # camera.h:41-46: n[6], p[6], corners[2][4] of vec3f
    assert capi.SEGMENT_DTYPE.itemsize == C.sizeof(capi.Segment) == 24
    assert C.sizeof(capi.HostArrays) == 11 * 8
    assert capi._ARRAY_SPEC[capi.BOUNDING_VOLUMES][1] == (13,)   # 52 B record, ecs_scene.h:248-255
    assert capi._ARRAY_SPEC[capi.POS_EXTENT][1] == (8,)          # 32 B record, ecs_scene.h:257-261


def test_no_device_means_loud_failure_not_a_fallback(lib):
    from pmtech_b200 import capi
    if capi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    from pmtech_b200.scene import DeviceScene
    with pytest.raises(capi.PmbError) as ei:
        DeviceScene(16)
    assert ei.value.code == -3  # PMB_ERR_NO_DEVICE
    assert "no CPU path" in str(ei.value)


def test_product_package_never_touches_the_oracle():
    """the oracle is test infrastructure: nothing under pmtech_b200/ may import, link or name it"""
    pkg = os.path.join(ROOT, "pmtech_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "pm_oracle" not in text and "from oracle" not in text and "import oracle" not in text, f


def _cam_pairs(oracle):
    from pmtech_b200.ecs import Camera
    return Camera, oracle.OracleCamera


def test_view_producers_match_the_oracle_bit_for_bit(lib, oracle):
    """camera_create_perspective / _orthographic / cubemap faces / look_at / update_frustum /
    update_shadow_frustum (core/put/source/camera.cpp:19-57,131-172,236-262,321-360,384-416)"""
    from pmtech_b200 import synth
    A, B = _cam_pairs(oracle)
    for mk in (synth.torus_views, synth.city_views, synth.mixed_views):
        va, vb = mk(A), mk(B)
        assert va.shape == vb.shape and va.shape[1] == 60
        assert np.array_equal(va.view(np.uint32), vb.view(np.uint32)), mk.__name__
    a = A().create_perspective(47.0, 1.3, 0.5, 250.0).look_at((3, 4, 5), (-1, 0.5, 2))
    b = B().create_perspective(47.0, 1.3, 0.5, 250.0).look_at((3, 4, 5), (-1, 0.5, 2))
    assert np.array_equal(a.view.view(np.uint32), b.view.view(np.uint32))
    assert np.array_equal(a.proj.view(np.uint32), b.proj.view(np.uint32))
    a = A().create_orthographic(-10, 12, -3, 4, 0.1, 90.0)
    b = B().create_orthographic(-10, 12, -3, 4, 0.1, 90.0)
    assert np.array_equal(a.proj.view(np.uint32), b.proj.view(np.uint32))


def test_frustum_planes_are_consistent(lib):
    """update_frustum: 6 planes from the 8 unprojected corners, order left, top, right, bottom,
    near, far (camera.cpp:152-161), outward normals; the 8 corners lie on their planes"""
    from pmtech_b200.ecs import Camera
    cam = Camera().create_perspective(60.0, 16.0 / 9.0, 0.1, 1000.0)
    cam.look_at((10, 20, 30), (0, 0, 0)).update_frustum()
    fr = cam.frustum
    n, p, corners = fr[0:18].reshape(6, 3), fr[18:36].reshape(6, 3), fr[36:60].reshape(8, 3)
    centre = corners.mean(axis=0)
    for k in range(6):
        nn = n[k] / np.linalg.norm(n[k])
        assert np.dot(centre - p[k], nn) < 0
        d = (corners - p[k]) @ nn
        assert (d < 1e-2 * max(1.0, np.abs(corners).max())).all()
        assert (np.abs(d) < 0.5).sum() >= 4  # four corners on each plane


def test_synthetic_scenes_respect_the_reference_invariants():
    from pmtech_b200 import synth
    for sc in (synth.torus_scene(), synth.deep_scene(n=1 << 16, block=8192), synth.city_scene(50_000),
               synth.rig_scene(100), synth.mixed_scene(4000)):
        n = sc["num_entities"]
        assert (sc["parents"] <= np.arange(n)).all()
        assert sc["local_matrices"].shape == (n, 16) and sc["local_matrices"].dtype == np.float32
        assert sc["bounding_volumes"].shape == (n, 13)
        assert sc["entities"].dtype == np.uint64 and sc["state_flags"].dtype == np.uint64
    assert synth.torus_scene()["num_entities"] == 65_544
    d = synth.depths_of(synth.deep_scene(n=1 << 16, block=8192)["parents"])
    assert d.max() == 16
    rig = synth.rig_scene(10)
    assert rig["num_entities"] == 660
    assert int(((rig["entities"] & np.uint64(synth.CMP_BONE)) != 0).sum()) == 640


def test_partition_by_root_and_slices():
    from pmtech_b200 import synth
    sc = synth.city_scene(200_000, seed=5)
    for parts in (2, 3, 8):
        ranges = synth.partition_by_root(sc["parents"], parts)
        assert ranges[0][0] == 0 and ranges[-1][1] == sc["num_entities"]
        for (a, b), (c, d) in zip(ranges[:-1], ranges[1:]):
            assert b == c
        for a, b in ranges:
            assert sc["parents"][a] == a                       # starts at a root
            assert (sc["parents"][a:b] >= a).all()             # closed under parent links
            sub = synth.slice_scene(sc, a, b)
            assert (sub["parents"] <= np.arange(b - a)).all()
            assert abs((b - a) - sc["num_entities"] / parts) < 0.05 * sc["num_entities"]


def test_instance_block_walk_matches_the_reference_loop(lib):
    """pmb_instance_blocks = the loop of ecs_scene.cpp:1994-2011 over host flags: masters without a custom buffer, their
    block = draw_call_data[n + 1 ..] of num_instances * instance_stride bytes, the walk striding over the sub-instances
    (a master flag INSIDE a block is therefore never seen)."""
    import ctypes as C
    from pmtech_b200 import capi
    n = 200
    MASTER, CUSTOM, SUB = 1 << 15, 1 << 21, 1 << 14
    ent = np.full(n, 1, np.uint64)
    mi = np.zeros(n, capi.MASTER_INSTANCE_DTYPE)
    layout = {3: (10, 41, 160), 14: (0, 42, 160), 20: (5, 43, 64), 40: (7, 44, 160), 199: (3, 45, 160)}
    for m, (k, buf, stride) in layout.items():
        ent[m] |= np.uint64(MASTER)
        ent[m + 1:m + 1 + k] |= np.uint64(SUB)
        mi[m] = (k, buf, stride)
    ent[40] |= np.uint64(CUSTOM)        # custom_instance_buffer: skipped, and its sub-instances are walked one by one
    ent[8] |= np.uint64(MASTER)         # inside master 3's block: strided over
    mi[8] = (2, 99, 160)
    want = []
    i = 0
    while i < n:                        # the reference loop, restated
        if (int(ent[i]) & MASTER) and not (int(ent[i]) & CUSTOM):
            want.append((i, int(mi[i]["instance_buffer"]), (i + 1) * 160, int(mi[i]["num_instances"]) * int(mi[i]["instance_stride"])))
            i += int(mi[i]["num_instances"])
        i += 1
    out = np.zeros(16, capi.INSTANCE_BLOCK_DTYPE)
    cnt = lib.pmb_instance_blocks(capi._ptr(ent), capi._ptr(mi), n, capi._ptr(out), 16)
    assert cnt == len(want) == 4
    got = [(int(b["master"]), int(b["instance_buffer"]), int(b["byte_offset"]), int(b["bytes"])) for b in out[:cnt]]
    assert got == want and [g[0] for g in got] == [3, 14, 20, 199]
    assert lib.pmb_instance_blocks(capi._ptr(ent), capi._ptr(mi), n, None, 0) == 4      # counting pass
    assert lib.pmb_instance_blocks(None, capi._ptr(mi), n, None, 0) == -1
