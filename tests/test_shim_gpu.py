"""The reference-side binding (include/put_ecs_b200.hpp) driven the way pmtech's call sites drive
the originals: ecs::update -> update_scene (ecs_scene.cpp:1322-1339), render_scene_view ->
filter_entities_scalar + frustum_cull_aabb + sb_count / sb_free (ecs_scene.cpp:957-960,1085-1093),
on a mock ecs_scene (tests/shim/mock_pmtech.h) whose cmp_arrays point at numpy buffers.  Results
are compared with the oracle bit for bit."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from helpers import run_oracle, ulp_diff

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM_DIR = os.path.join(ROOT, "tests", "shim")
SHIM_SO = os.path.join(SHIM_DIR, "libshim_driver.so")


def build_shim():
    from pmtech_b200 import build, capi
    if not os.path.exists(capi.SO_PATH):
        build.build()
    src = os.path.join(SHIM_DIR, "shim_driver.cpp")
    deps = [src, os.path.join(SHIM_DIR, "mock_pmtech.h"), os.path.join(ROOT, "include", "put_ecs_b200.hpp"),
            os.path.join(ROOT, "include", "pmtech_b200.h")]
    if not os.path.exists(SHIM_SO) or os.path.getmtime(SHIM_SO) < max(os.path.getmtime(d) for d in deps):
        libdir = os.path.dirname(capi.SO_PATH)
        subprocess.check_call(["g++", "-O2", "-std=c++11", "-fPIC", "-shared", "-Wall", "-I" + SHIM_DIR, "-o", SHIM_SO, src,
                               "-L" + libdir, "-lpmtech_b200", "-Wl,-rpath," + libdir])
    return SHIM_SO


def test_shim_compiles_against_the_mock_types():
    """no GPU needed: the header is C++11, builds with -Wall and links against the C ABI"""
    so = build_shim()
    L = C.CDLL(so)
    for s in ("shim_bind", "shim_update", "shim_update_opts", "shim_prepare_views", "shim_view_list", "shim_render_scene_view", "shim_cull_list",
              "shim_dir_shadow_view", "shim_release"):
        assert hasattr(L, s)


@pytest.mark.gpu
def test_shim_update_and_render_scene_view(gpu, oracle):
    from pmtech_b200 import synth
    from pmtech_b200.ecs import Camera
    L = C.CDLL(build_shim())
    L.shim_render_scene_view.restype = C.c_int
    L.shim_cull_list.restype = C.c_int
    sc = synth.mixed_scene(n=20000, seed=77)
    views = synth.mixed_views(Camera)
    n = sc["num_entities"]
    soa = n + 345  # capacity above num_entities, like a real scene
    arr = {
        "entities": np.zeros(soa, np.uint64), "state_flags": np.zeros(soa, np.uint64), "parents": np.zeros(soa, np.uint32),
        "local": np.zeros((soa, 16), np.float32), "world": np.zeros((soa, 16), np.float32),
        "bv": np.zeros((soa, 13), np.float32), "pe": np.zeros((soa, 8), np.float32),
    }
    arr["entities"][:n], arr["state_flags"][:n], arr["parents"][:n] = sc["entities"], sc["state_flags"], sc["parents"]
    arr["local"][:n], arr["world"][:n] = sc["local_matrices"], sc["world_matrices"]
    arr["bv"][:n], arr["pe"][:n] = sc["bounding_volumes"], sc["pos_extent"]
    # the mixed scene marks some physics entities as having an rb matrix; the shim has no physics
    # system to ask, so the oracle runs with "no rb matrices" too
    sc = dict(sc, has_rb_matrix=None)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    L.shim_bind(C.c_uint32(soa), C.c_uint32(n), p(arr["entities"]), p(arr["state_flags"]), p(arr["parents"]), p(arr["local"]),
                p(arr["world"]), p(arr["bv"]), p(arr["pe"]))
    ext = np.zeros(6, np.float32)
    L.shim_update(p(ext))
    want = run_oracle(oracle, sc, views)
    assert ulp_diff(arr["world"][:n], want["world"]) == 0
    assert ulp_diff(arr["bv"][:n], want["bv"]) == 0
    assert ulp_diff(arr["pe"][:n], want["pos_extent"]) == 0
    assert ulp_diff(ext, want["extents"]) == 0
    assert np.array_equal(arr["entities"][:n], want["entities"])
    assert np.array_equal(arr["state_flags"][:n], want["state_flags"])
    assert not arr["world"][n:].any()  # beyond num_entities nothing is touched
    out = np.zeros(n, np.uint32)
    for v, fr in enumerate(views):
        fr = np.ascontiguousarray(fr, np.float32)
        cnt = L.shim_render_scene_view(p(fr), 0, p(out), n)
        assert cnt == len(want["lists"][v])
        assert np.array_equal(out[:cnt], want["lists"][v]), v
    # sphere variant
    want_s = run_oracle(oracle, sc, views, mode="sphere")
    fr = np.ascontiguousarray(views[1], np.float32)
    cnt = L.shim_render_scene_view(p(fr), 1, p(out), n)
    assert np.array_equal(out[:cnt], want_s["lists"][1])
    # a caller-ordered sub-list keeps its order (ecs_cull.cpp:29-49)
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc["world_matrices"], pos_extent=sc["pos_extent"])
    o.update_scene()
    sub = o.filter_entities()[::3][::-1].copy()
    want_sub = o.frustum_cull(views[0], sub, "aabb")
    cnt = L.shim_cull_list(p(np.ascontiguousarray(views[0], np.float32)), p(sub), len(sub), p(out))
    assert np.array_equal(out[:cnt], want_sub)
    # a directional light's shadow view, fitted on the device from the extents the update left there
    from test_shadow_fit_gpu import _fit
    L.shim_dir_shadow_view.restype = C.c_int
    direction = np.array([0.3, -1.0, 0.2], np.float32)
    cam_rec = np.zeros(32 + 60, np.float32)
    cnt = L.shim_dir_shadow_view(p(direction), p(out), n, p(cam_rec))
    cam = _fit(oracle, want["extents"], direction.reshape(1, 3))[0]
    assert np.array_equal(cam_rec[:16].view(np.uint32), cam.view.view(np.uint32))
    assert np.array_equal(cam_rec[16:32].view(np.uint32), cam.proj.view(np.uint32))
    assert np.array_equal(cam_rec[32:].view(np.uint32), cam.frustum.view(np.uint32))
    want_shadow = o.frustum_cull(cam.frustum, o.filter_entities(), "aabb")
    assert cnt == len(want_shadow) > 0 and np.array_equal(out[:cnt], want_shadow)
    L.shim_release()


def _bind(L, arr, soa, n):
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    L.shim_bind(C.c_uint32(soa), C.c_uint32(n), p(arr["entities"]), p(arr["state_flags"]), p(arr["parents"]), p(arr["local"]),
                p(arr["world"]), p(arr["bv"]), p(arr["pe"]))


def _host_arrays(sc, soa):
    n = sc["num_entities"]
    arr = {"entities": np.zeros(soa, np.uint64), "state_flags": np.zeros(soa, np.uint64), "parents": np.zeros(soa, np.uint32),
           "local": np.zeros((soa, 16), np.float32), "world": np.zeros((soa, 16), np.float32),
           "bv": np.zeros((soa, 13), np.float32), "pe": np.zeros((soa, 8), np.float32)}
    arr["entities"][:n], arr["state_flags"][:n], arr["parents"][:n] = sc["entities"], sc["state_flags"], sc["parents"]
    arr["local"][:n] = sc["local_matrices"]
    arr["bv"][:n] = sc["bounding_volumes"]
    if sc.get("world_matrices") is not None:
        arr["world"][:n], arr["pe"][:n] = sc["world_matrices"], sc["pos_extent"]
    return arr


@pytest.mark.gpu
def test_shim_resident_form_dirty_ranges_rb_matrices_and_the_view_cache(gpu, oracle):
    """The binding a pmtech build would ship (frame_options): frame 1 sends everything and honours physics::has_rb_matrix;
    frame 2 sends only two dirtied ranges of local matrices + flags, reads nothing back but the extents, clears the
    host's dirty flags itself, and serves the six render_scene_view calls from ONE batched device pass; frame 3 reads
    the outputs back.  Every frame equals the oracle's."""
    from pmtech_b200 import synth
    from pmtech_b200.ecs import Camera
    L = C.CDLL(build_shim())
    L.shim_render_scene_view.restype = C.c_int
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    sc = synth.mixed_scene(n=20000, seed=78)
    views = synth.mixed_views(Camera)
    n = sc["num_entities"]
    soa = n + 99
    arr = _host_arrays(sc, soa)
    has_rb = np.zeros(soa, np.uint8)
    has_rb[:n] = sc["has_rb_matrix"]
    assert has_rb.any()
    _bind(L, arr, soa, n)
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc["world_matrices"], pos_extent=sc["pos_extent"], has_rb_matrix=sc["has_rb_matrix"])
    ext = np.zeros(6, np.float32)
    # frame 1: everything up, everything down
    L.shim_update_opts(p(ext), p(has_rb), 1, None, C.c_uint32(0), None, 0, 15)
    o.update_scene()
    assert ulp_diff(arr["world"][:n], o.world) == 0 and ulp_diff(arr["bv"][:n], o.bv) == 0 and ulp_diff(arr["pe"][:n], o.pos_extent) == 0
    assert np.array_equal(arr["entities"][:n], o.entities) and np.array_equal(arr["state_flags"][:n], o.state_flags)
    assert ulp_diff(ext, o.renderable_extents) == 0
    world1 = arr["world"].copy()
    # frame 2: the host moves two ranges of entities and flags them (what every pmtech site that edits a transform does)
    rng = np.random.default_rng(9)
    ranges = np.array([[100, 500], [7000, 1]], np.uint32)
    for first, count in ranges:
        sl = slice(int(first), int(first + count))
        move = rng.uniform(-20.0, 20.0, (int(count), 3)).astype(np.float32)
        for a in (arr["local"], o.local):
            a[sl, 3] += move[:, 0]
            a[sl, 7] += move[:, 1]
            a[sl, 11] += move[:, 2]
        arr["entities"][sl] |= np.uint64(1 << 12)
        o.entities[sl] |= np.uint64(1 << 12)
    L.shim_update_opts(p(ext), p(has_rb), 0, p(ranges), C.c_uint32(len(ranges)), None, 0, 0)
    o.update_scene()
    assert ulp_diff(ext, o.renderable_extents) == 0
    assert np.array_equal(arr["entities"][:n], o.entities) and np.array_equal(arr["state_flags"][:n], o.state_flags)
    assert np.array_equal(arr["world"], world1)                     # nothing came down
    assert ulp_diff(world1[:n], o.world) > 0                        # ... although the frame did change the scene
    L.shim_prepare_views(p(np.ascontiguousarray(views, np.float32)), C.c_uint32(len(views)), 0)
    filt = o.filter_entities()
    out = np.zeros(n, np.uint32)
    for v, fr in enumerate(views):
        want = o.frustum_cull(fr, filt, "aabb")
        cnt = L.shim_render_scene_view(p(np.ascontiguousarray(fr, np.float32)), 0, p(out), n)
        assert cnt == len(want) and np.array_equal(out[:cnt], want), v
    L.shim_view_list.restype = C.c_int
    for v, fr in enumerate(views):  # the pair as one call
        want = o.frustum_cull(fr, filt, "aabb")
        cnt = L.shim_view_list(p(np.ascontiguousarray(fr, np.float32)), 0, p(out), n)
        assert cnt == len(want) and np.array_equal(out[:cnt], want), v
    # a view that was not prepared is culled on demand and cached too
    other = np.ascontiguousarray(synth.torus_views(Camera)[0], np.float32)
    want = o.frustum_cull(other, filt, "aabb")
    for _ in range(2):
        cnt = L.shim_render_scene_view(p(other), 0, p(out), n)
        assert cnt == len(want) and np.array_equal(out[:cnt], want)
    # frame 3: nothing dirtied, outputs read back
    L.shim_update_opts(p(ext), p(has_rb), 0, None, C.c_uint32(0), None, 0, 7)
    o.update_scene()
    assert ulp_diff(arr["world"][:n], o.world) == 0 and ulp_diff(arr["bv"][:n], o.bv) == 0 and ulp_diff(arr["pe"][:n], o.pos_extent) == 0
    L.shim_release()


@pytest.mark.gpu
def test_shim_resident_form_with_stage_0_on_the_device(gpu, oracle):
    """cmp_transform records (40 B) travel instead of local matrices, e_cmp::transform is set for the dirtied range on the
    device, and local = T * R * S is baked there (ecs_scene.cpp:1415-1427): world matrices equal the oracle's, which bakes too."""
    from pmtech_b200 import synth
    L = C.CDLL(build_shim())
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    sc = synth.city_scene(30000, seed=12, with_transforms=True)
    n = sc["num_entities"]
    soa = n
    sc["entities"] = sc["entities"] | np.uint64(1 << 12)   # every transform dirtied by the host
    arr = _host_arrays(sc, soa)
    tr = np.ascontiguousarray(sc["transforms"], np.float32)
    _bind(L, arr, soa, n)
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           transforms=tr)
    ext = np.zeros(6, np.float32)
    L.shim_update_opts(p(ext), None, 1, None, C.c_uint32(0), p(tr), 0, 15)
    o.update_scene()
    assert ulp_diff(arr["world"][:n], o.world) == 0 and ulp_diff(ext, o.renderable_extents) == 0
    assert np.array_equal(arr["entities"][:n], o.entities)
    # frame 2: a range of transforms moves; only those 40 B records go up, the flag is set on the device.
    # (get_matrix normalises the rotations in place, ecs_scene.cpp:1421: the device did that to its mirror, the oracle to
    # its array; the host array of this test takes the oracle's, as pmtech's own array would hold them)
    tr[:] = o.transforms
    ranges = np.array([[5000, 3000]], np.uint32)
    tr[5000:8000, 0] += np.float32(1.5)
    o.transforms[5000:8000, 0] += np.float32(1.5)
    o.entities[5000:8000] |= np.uint64(1 << 12)
    L.shim_update_opts(p(ext), None, 0, p(ranges), C.c_uint32(1), p(tr), 1, 1)
    o.update_scene()
    assert ulp_diff(arr["world"][:n], o.world) == 0 and ulp_diff(ext, o.renderable_extents) == 0
    assert (arr["world"][5000:8000, 3] != sc["local_matrices"][5000:8000, 3]).any()
    L.shim_release()
