"""Shadow views fitted on the device: pmb_cull_shadow_views against the host producers.

The reference builds a directional light's shadow camera on the host from scene->renderable_extents
(shadow_camera_from_entity, ecs_scene.cpp:538-554 -> camera_update_shadow_frustum, camera.cpp:384-416) and
then culls the scene with its frustum.  The device call must give the cameras the oracle's restatement of
those producers gives for the oracle's extents — view, proj and all 60 frustum floats bit for bit — and the
draw lists the oracle's cull gives for those frusta."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

F = np.float32


def _light_dir(d):
    """normalize(-direction) in binary32, one operation at a time (dot left to right, then three divisions)"""
    v = -np.asarray(d, F)
    m = np.sqrt(F(F(F(v[0] * v[0]) + F(v[1] * v[1])) + F(v[2] * v[2])))
    return np.array([v[0] / m, v[1] / m, v[2] / m], F)


def _fit(orc, extents, directions, constraints=None):
    emin, emax = np.asarray(extents[:3], F).copy(), np.asarray(extents[3:], F).copy()
    if constraints is not None:
        c = np.asarray(constraints, F)
        d = c[:3] - c[3:]
        if F(F(F(d[0] * d[0]) + F(d[1] * d[1])) + F(d[2] * d[2])) != 0:
            emin, emax = np.maximum(c[:3], emin), np.minimum(c[3:], emax)
    cams = []
    for d in directions:
        cam = orc.OracleCamera()
        cam.update_shadow_frustum(_light_dir(d), emin - F(0.1), emax + F(0.1))
        cams.append(cam)
    return cams


def _bits(a):
    return np.ascontiguousarray(a, F).view(np.uint32)


DIRECTIONS = np.array([[0.3, -1.0, 0.2], [-0.5, -0.7, 0.5], [0.9, -0.1, -0.4]], F)


@pytest.mark.parametrize("mode", ["aabb", "sphere"])
@pytest.mark.parametrize("constraints", [None, (-40, -5, -30, 35, 60, 45), (1, 2, 3, 1, 2, 3)])
def test_device_shadow_views_match_the_host_producers(oracle, mode, constraints):
    from pmtech_b200 import synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.mixed_scene(n=3000, seed=21)
    n = sc["num_entities"]
    views = synth.mixed_views(oracle.OracleCamera)
    nv = len(views)
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc.get("world_matrices"), pos_extent=sc.get("pos_extent"),
                           has_rb_matrix=sc.get("has_rb_matrix"))
    o.update_scene()
    ext = o.renderable_extents
    ext = np.asarray(ext() if callable(ext) else ext, F)
    cams = _fit(oracle, ext, DIRECTIONS, constraints)
    shadow_frusta = np.stack([c.frustum for c in cams])
    want, _ = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                                 world_matrices=sc.get("world_matrices"), pos_extent=sc.get("pos_extent"),
                                 has_rb_matrix=sc.get("has_rb_matrix")).frame(
        np.concatenate([np.asarray(views, F).reshape(nv, 60), shadow_frusta]), mode=mode)
    with DeviceScene(n, 0) as dev:
        upload_scene(dev, sc)
        dev.reserve_views(nv + 3)
        dev.frame(n, views, mode)
        dev.cull_shadow_views(n, DIRECTIONS, constraints, mode, first_view=nv)
        got = dev.lists(nv + 3)
        rec = dev.download_shadow_cameras(3)
        assert np.array_equal(_bits(dev.renderable_extents()), _bits(ext))
    for l, cam in enumerate(cams):
        assert np.array_equal(_bits(rec[l, 3:19]), _bits(cam.view)), ("view", l)
        assert np.array_equal(_bits(rec[l, 19:35]), _bits(cam.proj)), ("proj", l)
        assert np.array_equal(_bits(rec[l, 35:95]), _bits(cam.frustum)), ("frustum", l)
    assert len(got) == len(want) == nv + 3
    for v in range(nv + 3):
        assert np.array_equal(got[v], want[v]), v
    sizes = [len(want[nv + l]) for l in range(3)]
    assert max(sizes) > 0


def test_shadow_views_alone_and_the_host_wrapper_agree(oracle):
    """first_view = 0 needs no reservation; the product's own host producer gives the same camera"""
    from pmtech_b200 import ecs, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.mixed_scene(n=2000, seed=5)
    n = sc["num_entities"]
    with DeviceScene(n, 0) as dev:
        upload_scene(dev, sc)
        dev.update_scene(n)
        dev.cull_shadow_views(n, DIRECTIONS[:2])
        lists = dev.lists(2)
        rec = dev.download_shadow_cameras(2)
        ext = dev.renderable_extents()
        for l in range(2):
            cam = ecs.Camera().update_shadow_frustum(_light_dir(DIRECTIONS[l]), ext[:3] - F(0.1), ext[3:] + F(0.1))
            assert np.array_equal(_bits(rec[l, 35:95]), _bits(cam.frustum))
            dev.cull_views(n, cam.frustum.reshape(1, 60), "aabb")
            assert np.array_equal(dev.lists(1)[0], lists[l])
        assert len(lists[0]) > 0 and len(lists[1]) > 0


def test_unreserved_append_is_refused():
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.mixed_scene(n=500, seed=1)
    with DeviceScene(500, 0) as dev:
        upload_scene(dev, sc)
        dev.update_scene(500)
        with pytest.raises(capi.PmbError) as e:
            dev.cull_shadow_views(500, DIRECTIONS, first_view=40)
        assert "reserve" in str(e.value)
        with pytest.raises(capi.PmbError):
            dev.cull_shadow_views(500, np.zeros((17, 3), F))


@pytest.mark.parametrize("mode", ["aabb", "sphere"])
def test_device_spot_shadow_views_match_the_host_producers(oracle, mode):
    """the spot branch of shadow_camera_from_entity (ecs_scene.cpp:556-570): cameras from the lights' world matrices,
    which only exist on the device after the update (children of moving parents included)"""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.mixed_scene(n=3000, seed=33)
    n = sc["num_entities"]
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc.get("world_matrices"), pos_extent=sc.get("pos_extent"),
                           has_rb_matrix=sc.get("has_rb_matrix"))
    o.update_scene()
    # "lights": entities with a well-conditioned world matrix, at least one of them a child
    det = np.abs(np.linalg.det(o.world.reshape(n, 4, 4)[:, :3, :3].astype(np.float64)))
    child = sc["parents"] != np.arange(n)
    ok = np.flatnonzero((det > 1e-3) & np.isfinite(o.world).all(axis=1))
    lights = np.concatenate([ok[~child[ok]][:2], ok[child[ok]][:2]]).astype(np.uint32)
    assert len(lights) == 4 and child[lights].any()
    cams = [oracle.OracleCamera().spot_shadow(o.world[e]) for e in lights]
    frusta = np.stack([c.frustum for c in cams])
    want, _ = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                                 world_matrices=sc.get("world_matrices"), pos_extent=sc.get("pos_extent"),
                                 has_rb_matrix=sc.get("has_rb_matrix")).frame(frusta, mode=mode)
    with DeviceScene(n, 0) as dev:
        upload_scene(dev, sc)
        dev.update_scene(n)
        dev.cull_spot_shadow_views(n, lights, mode)
        got = dev.lists(4)
        rec = dev.download_shadow_cameras(4)
        with pytest.raises(capi.PmbError):
            dev.cull_spot_shadow_views(n, [n + 5], mode)
    for l, cam in enumerate(cams):
        assert np.array_equal(_bits(rec[l, 3:19]), _bits(cam.view)), ("view", l)
        assert np.array_equal(_bits(rec[l, 19:35]), _bits(cam.proj)), ("proj", l)
        assert np.array_equal(_bits(rec[l, 35:95]), _bits(cam.frustum)), ("frustum", l)
        assert np.array_equal(got[l], want[l]), l
    assert sum(len(w) for w in want) > 0
