"""Parity of the CUDA path (through the C ABI) against the oracle on the same seeded scenes:
every BASELINE.json config at a size the oracle finishes in seconds, plus the edge cases the
reference loops carry (SURVEY.md §3.2-3.4, §7 hard part 6)."""
import numpy as np
import pytest

from helpers import assert_parity, run_device, run_oracle, ulp_diff, upload_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cams(gpu):
    from pmtech_b200.ecs import Camera
    return Camera


def test_config1_torus_64k(gpu, oracle, cams):
    from pmtech_b200 import synth
    sc = synth.torus_scene()
    views = synth.torus_views(cams)
    want = run_oracle(oracle, sc, views)
    got = run_device(sc, views)
    worst = assert_parity(got, want, "torus")
    assert 0 < len(got["lists"][0]) < sc["num_entities"]
    assert max(worst.values()) == 0, worst  # canonical mode: same operation order, no contraction


def test_config2_deep_hierarchy_1m(gpu, oracle):
    from pmtech_b200 import synth
    sc = synth.deep_scene()
    assert sc["depth"].max() == 16
    want = run_oracle(oracle, sc, None)
    got = run_device(sc, None)
    worst = assert_parity(got, want, "deep 1M")
    assert max(worst.values()) == 0, worst


@pytest.mark.parametrize("fused", [True, False])
def test_config3_city_11_views_1m(gpu, oracle, cams, fused):
    from pmtech_b200 import synth
    sc = synth.city_scene(1 << 20)
    views = synth.city_views(cams)
    assert views.shape == (11, 60)
    want = run_oracle(oracle, sc, views)
    got = run_device(sc, views, fused=fused)
    worst = assert_parity(got, want, f"city 1M fused={fused}")
    assert max(worst.values()) == 0, worst
    assert sum(len(l) for l in got["lists"]) > 0


def test_config4_rigs(gpu, oracle, cams):
    from pmtech_b200 import synth
    sc = synth.rig_scene(10_000)  # BASELINE config 4 at its full size
    views = synth.mixed_views(cams)
    want = run_oracle(oracle, sc, views)
    got = run_device(sc, views)
    worst = assert_parity(got, want, "rigs")
    assert max(worst.values()) == 0, worst


@pytest.mark.parametrize("mode", ["aabb", "sphere"])
@pytest.mark.parametrize("n", [1, 31, 255, 256, 257, 1023, 1025, 5000, 40001])
def test_mixed_edge_cases(gpu, oracle, cams, n, mode):
    """ragged sizes around the CTA / compaction tiles; chains deeper than the in-thread walk (the
    wavefront form of loop A); unallocated slots; bones; hidden / sub_instance / no-material; physics with and
    without an rb matrix; dirty flags cleared."""
    from pmtech_b200 import synth
    sc = synth.mixed_scene(n=n, seed=100 + n)
    views = synth.mixed_views(cams)
    want = run_oracle(oracle, sc, views, mode=mode)
    got = run_device(sc, views, mode=mode)
    worst = assert_parity(got, want, f"mixed n={n} {mode}")
    assert max(worst.values()) == 0, worst
    got2 = run_device(sc, views, mode=mode, fused=False, capacity=n + 77)
    assert_parity(got2, want, f"mixed two-call n={n} {mode}")


def test_empty_scene_and_no_views(gpu, oracle, cams):
    from pmtech_b200 import synth
    from pmtech_b200.scene import DeviceScene
    views = synth.mixed_views(cams)
    with DeviceScene(16) as dev:
        dev.frame(0, views)
        assert dev.counts(len(views)).tolist() == [0] * len(views)
        assert all(d.size == 0 and h is None for d, h in dev.lists_sb(len(views)))
        ext = dev.renderable_extents()
        f = np.finfo(np.float32).max
        assert ext.tolist() == [f, f, f, -f, -f, -f]  # ecs_scene.cpp:1482-1483 with nothing folded
        dev.frame(0, np.zeros((0, 60), np.float32))


def test_more_than_16_views_and_filter_only(gpu, oracle, cams):
    from pmtech_b200 import synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.mixed_scene(n=20000, seed=5)
    base = synth.mixed_views(cams)
    views = np.concatenate([base] * 6)[:35]  # 35 views: three passes of the 16-view batch
    want = run_oracle(oracle, sc, views)
    got = run_device(sc, views)
    assert_parity(got, want, "35 views")
    o = run_oracle(oracle, sc, None)
    filt = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"],
                              sc["bounding_volumes"]).filter_entities()
    with DeviceScene(sc["num_entities"]) as dev:
        upload_scene(dev, sc)
        dev.cull_views(sc["num_entities"], base[:1], "filter")
        assert np.array_equal(dev.lists(1)[0], filt)
    del o


def test_stretchy_buffer_lists(gpu, oracle, cams):
    """pmb_get_lists_sb returns buffers the reference's sb_count / sb_free accept
    (core/pen/include/data_struct.h:22-72): {int cap, int count} header 8 bytes before the data."""
    from pmtech_b200 import synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.torus_scene()
    views = synth.torus_views(cams)
    want = run_oracle(oracle, sc, views)
    with DeviceScene(sc["num_entities"]) as dev:
        upload_scene(dev, sc)
        dev.frame(sc["num_entities"], views)
        (data, hdr), = dev.lists_sb(1)
        assert hdr[1] == len(want["lists"][0]) and hdr[0] >= hdr[1]
        assert np.array_equal(data, want["lists"][0])


def test_second_frame_is_idempotent_and_flags_stay_clear(gpu, oracle, cams):
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import download_outputs, upload_scene
    sc = synth.mixed_scene(n=30000, seed=11)
    views = synth.mixed_views(cams)
    n = sc["num_entities"]
    with DeviceScene(n) as dev:
        upload_scene(dev, sc)
        dev.frame(n, views)
        a = download_outputs(dev, n, len(views))
        dev.frame(n, views)
        b = download_outputs(dev, n, len(views))
    # frame 2 on the oracle = the oracle run on frame 1's outputs
    sc2 = dict(sc, entities=a["entities"], state_flags=a["state_flags"], world_matrices=a["world"],
               pos_extent=a["pos_extent"], bounding_volumes=a["bv"])
    want2 = run_oracle(oracle, sc2, views)
    assert_parity(b, want2, "second frame")
    assert not (b["entities"] & np.uint64(synth.CMP_TRANSFORM)).any()


def test_one_launch_culls_between_updates_and_on_a_scene_that_shrinks_and_grows(gpu, oracle, cams):
    """The one-launch frame keeps two sets of group / chunk totals and each cull launch clears the other set.  An
    update-only launch in between (pmb_update_scene, pmb_frame without views) must not change which set the next cull
    uses, and a frame on fewer entities must not leave totals of the larger frame behind: every cull of the sequence
    gives the oracle's lists (doubled counts / shifted tile offsets otherwise)."""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    sc = synth.city_scene(200_000)  # 3125 tiles: 98 groups, 4 chunks
    views = synth.city_views(cams)
    n, part = sc["num_entities"], 70_000  # 1094 tiles: 35 groups, 2 chunks
    want = {n: run_oracle(oracle, sc, views)["lists"], part: run_oracle(oracle, synth.slice_scene(sc, 0, part), views)["lists"]}
    assert sum(len(l) for l in want[part]) > 0
    with DeviceScene(n) as dev:
        dev.set_option(capi.OPT_FUSED_TAIL, 2)  # one launch whenever the scene is small enough
        for step, m in enumerate((n, 0, n, part, n, 0, 0, part, part, n, 0, n)):
            upload_scene(dev, sc)  # every step starts from the same inputs
            if m == 0:
                if step % 2:
                    dev.update_scene(n)
                else:
                    dev.frame(n, views[:0])
                continue
            dev.frame(m, views)
            got = dev.lists(len(views))
            for v, (g, w) in enumerate(zip(got, want[m])):
                assert g.shape == w.shape, f"step {step} ({m} entities) view {v}: {g.shape[0]} visible, oracle {w.shape[0]}"
                assert np.array_equal(g, w), f"step {step} ({m} entities) view {v}: draw list differs"


def test_frame_host_round_trip(gpu, oracle, cams):
    """the host-buffer entry: uploads, frame, list / output read-back in one call"""
    from pmtech_b200 import synth
    from pmtech_b200.scene import DeviceScene
    sc = synth.city_scene(200_000, seed=9)
    views = synth.city_views(cams)
    n, nv = sc["num_entities"], len(views)
    want = run_oracle(oracle, sc, views)
    world = np.empty((n, 16), np.float32)
    pe = np.empty((n, 8), np.float32)
    bv = sc["bounding_volumes"].copy()
    ext = np.empty(6, np.float32)
    lists = [np.empty(n, np.uint32) for _ in range(nv)]
    with DeviceScene(n) as dev:
        cnt = dev.frame_host(n, views, parents=sc["parents"], local_matrices=sc["local_matrices"],
                             entities=sc["entities"], state_flags=sc["state_flags"],
                             bounding_volumes=sc["bounding_volumes"], world_matrices_out=world, pos_extent_out=pe,
                             bounding_volumes_out=bv, renderable_extents_out=ext, lists_out=lists)
    for v in range(nv):
        assert np.array_equal(lists[v][:cnt[v]], want["lists"][v])
    assert ulp_diff(world, want["world"]) == 0
    assert ulp_diff(pe, want["pos_extent"]) == 0
    assert ulp_diff(bv, want["bv"]) == 0
    assert ulp_diff(ext, want["extents"]) == 0


def test_resize_keeps_contents(gpu, oracle, cams):
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import download_outputs, upload_scene
    sc = synth.mixed_scene(n=3000, seed=21)
    views = synth.mixed_views(cams)
    want = run_oracle(oracle, sc, views)
    with DeviceScene(3000) as dev:
        upload_scene(dev, sc)
        dev.resize(10_000)  # resize_scene_buffers (ecs_scene.cpp:122-152)
        dev.frame(3000, views)
        got = download_outputs(dev, 3000, len(views))
        tail = dev.download(capi.WORLD_MATRICES, 3000, 7000)
    assert_parity(got, want, "after resize")
    assert not tail.any()


def test_errors_are_loud(gpu):
    from pmtech_b200 import capi
    from pmtech_b200.scene import DeviceScene
    with DeviceScene(10) as dev:
        with pytest.raises(capi.PmbError):
            dev.update_scene(11)  # beyond capacity
        with pytest.raises(capi.PmbError):
            dev.upload(capi.PARENTS, np.zeros(11, np.uint32))
        with pytest.raises(capi.PmbError):
            dev.counts(1)  # lists read before any cull


def test_full_size_16m_against_the_oracle(gpu, cams):
    """BASELINE config 3 at its full size (16,777,216 entities, 11 views): every draw list, every world matrix, every
    bounding volume and pos_extent compared bit for bit with the oracle (its OpenMP form, itself tested bit-identical
    to the serial one in test_oracle_cpu.py), plus the size-independent properties: lists strictly ascending and
    inside the filter set, the fused and the two-call forms agree, idempotent second frame."""
    from oracle import oracle as orc
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    n = synth.N_16M
    sc = synth.city_scene(n)
    views = synth.city_views(cams)
    nv = len(views)
    with DeviceScene(n) as dev:
        upload_scene(dev, sc)
        dev.frame(n, views)
        lists = dev.lists(nv)
        ext = dev.renderable_extents()
        dev.update_scene(n)
        dev.cull_views(n, views)
        lists2 = dev.lists(nv)
        ext2 = dev.renderable_extents()
        pe = dev.download(capi.POS_EXTENT, 0, n)
        bv = dev.download(capi.BOUNDING_VOLUMES, 0, n)
        world = dev.download(capi.WORLD_MATRICES, 0, n)
    renderable = ((sc["entities"] & np.uint64(0x12)) == np.uint64(0x12)) & ((sc["state_flags"] & np.uint64(4)) == 0)
    for v in range(nv):
        l = lists[v].astype(np.int64)
        assert np.array_equal(lists[v], lists2[v])
        assert (np.diff(l) > 0).all()
        assert renderable[l].all()
    assert np.array_equal(ext, ext2)
    assert (pe[:, 4:7] >= 0).all()
    want = run_oracle(orc, sc, views, omp=True)
    for v in range(nv):
        assert np.array_equal(lists[v], want["lists"][v]), f"view {v}"
    assert np.array_equal(world.view(np.uint32), np.asarray(want["world"]).view(np.uint32))
    assert np.array_equal(pe.view(np.uint32), np.asarray(want["pos_extent"]).view(np.uint32))
    assert np.array_equal(bv.view(np.uint32), np.asarray(want["bv"]).view(np.uint32))
    assert np.array_equal(ext.view(np.uint32), np.asarray(want["extents"], np.float32).view(np.uint32))


@pytest.mark.parametrize("name", ["mixed_3000", "torus_views", "rigs_40"])
def test_golden_fixtures_on_device(gpu, cams, name):
    """the CUDA path against the committed vectors of tests/golden/ (no oracle in the loop)"""
    import os
    import make_golden_lib as mg
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    sc, views, mode = mg.scene_for(name, cams)
    assert np.array_equal(views.view(np.uint32), z["views"].view(np.uint32))
    got = run_device(sc, views, mode=mode)
    for k in ("world", "bv", "pos_extent", "extents"):
        assert ulp_diff(got[k], z[k]) == 0, k
    assert np.array_equal(got["entities"], z["entities"])
    assert np.array_equal(got["state_flags"], z["state_flags"])
    for v in range(len(views)):
        assert np.array_equal(got["lists"][v], z[f"list{v}"]), v


@pytest.mark.gpu
@pytest.mark.parametrize("world", [1, 2, 3])
def test_row_exchange_kernels_match_the_host_statement(gpu, oracle, cams, world):
    """pmb_pack_row / pmb_unpack_rows (the multi-GPU list exchange, shard.py) on one GPU: the shards of
    a scene are culled one after the other, their rows land in one buffer as an all_gather would leave
    them, and the unpacked per-view lists must equal the unsharded scene's lists."""
    import torch
    from pmtech_b200 import shard, synth
    from pmtech_b200.scene import DeviceScene
    n = 50_000
    sc = synth.city_scene(n, seed=23)
    views = synth.city_views(cams)
    nv = len(views)
    want = run_oracle(oracle, sc, views)["lists"]
    parts = synth.partition_by_root(sc["parents"], world)
    stride = (max(sum(len(l) for l in want), 1) + nv + 7) & ~3
    rows = torch.zeros(world * stride, dtype=torch.int32, device="cuda")
    host_rows = []
    for r, (first, last) in enumerate(parts):
        sub = synth.slice_scene(sc, first, last)
        with DeviceScene(last - first, 0) as dev:
            upload_scene(dev, sub)
            dev.frame(last - first, views)
            dev.pack_row(nv, first, rows.data_ptr() + 4 * r * stride, stride)
            dev.synchronize()
            host_rows.append(shard.pack_row_np(dev.lists(nv), first, stride))
    assert np.array_equal(rows.cpu().numpy().view(np.uint32).reshape(world, stride)[:, :], np.stack(host_rows))
    out = torch.zeros(world * stride, dtype=torch.int32, device="cuda")
    meta = torch.zeros(2 * nv + 1, dtype=torch.int32, device="cuda")
    with DeviceScene(16, 0) as dev:
        dev.unpack_rows(rows.data_ptr(), world, stride, nv, out.data_ptr(), out.numel(), meta.data_ptr())
        dev.synchronize()
        m = meta.cpu().numpy()
        assert m[2 * nv] == 0
        o = out.cpu().numpy().view(np.uint32)
        for v in range(nv):
            assert np.array_equal(o[m[v]:m[v] + m[nv + v]], want[v]), v
        # a destination that is too small is reported and left alone
        meta.zero_()
        dev.unpack_rows(rows.data_ptr(), world, stride, nv, out.data_ptr(), max(int(m[nv:2 * nv].sum()) - 1, 0), meta.data_ptr())
        dev.synchronize()
        assert meta.cpu().numpy()[2 * nv] == 1 or int(m[nv:2 * nv].sum()) == 0


def test_config4_stage0_root_motion_and_skin_palettes(gpu, oracle, cams):
    """BASELINE config 4 end to end on the device: root-motion offsets land on the rig roots' transforms,
    stage 0 bakes local = T * R * S for every flagged entity, the frame propagates the 64-bone
    hierarchies, and the skin palettes world * bind come out — all bit for bit the oracle's."""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    sc = synth.rig_scene(1500, seed=4)
    n = sc["num_entities"]
    rng = np.random.default_rng(11)
    # the animation dirtied every joint; a few rigs are static this frame, a few entities sync with physics
    flagged = np.repeat(rng.random(1500) < 0.9, synth.RIG_ENTITIES)
    sc["entities"] = sc["entities"] | np.where(flagged, np.uint64(1 << 12), np.uint64(0))
    sc["state_flags"] = sc["state_flags"] | np.where(rng.random(n) < 0.01, np.uint64(1 << 7), np.uint64(0))
    views = synth.mixed_views(cams)
    rm = sc["root_motion"]

    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"],
                           transforms=sc["transforms"])
    o.apply_root_motion(rm["targets"], rm["rotations"], rm["deltas"])
    want_lists, _ = o.frame(views, mode="sphere")
    want_pal = o.skin_palettes(sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])

    with DeviceScene(n, 0) as dev:
        upload_scene(dev, sc)
        dev.set_skins(sc["skins"], sc["bind_matrices"], sc["num_palette_matrices"])
        dev.apply_root_motion(rm["targets"], rm["rotations"], rm["deltas"])
        dev.frame(n, views, "sphere")
        dev.skin_palettes()
        got_pal = dev.download_palettes()
        assert np.array_equal(dev.download(capi.TRANSFORMS, 0, n).view(np.uint32), o.transforms.view(np.uint32))
        assert np.array_equal(dev.download(capi.LOCAL_MATRICES, 0, n).view(np.uint32), o.local.view(np.uint32))
        assert np.array_equal(dev.download(capi.WORLD_MATRICES, 0, n).view(np.uint32), o.world.view(np.uint32))
        assert np.array_equal(dev.download(capi.BOUNDING_VOLUMES, 0, n).view(np.uint32), o.bv.view(np.uint32))
        assert np.array_equal(dev.download(capi.ENTITIES, 0, n), o.entities)
        assert np.array_equal(dev.download(capi.STATE_FLAGS, 0, n), o.state_flags)
        lists = dev.lists(len(views))
        for v in range(len(views)):
            assert np.array_equal(lists[v], want_lists[v]), v
        assert np.array_equal(got_pal.view(np.uint32), want_pal.view(np.uint32))
        assert sum(len(l) for l in lists) > 0
        # second frame: nothing is flagged any more, so nothing is re-baked and the results stand
        dev.frame(n, views, "sphere")
        assert np.array_equal(dev.download(capi.WORLD_MATRICES, 0, n).view(np.uint32), o.world.view(np.uint32))


def test_root_motion_needs_transforms_and_skins_are_checked(gpu):
    from pmtech_b200 import capi
    from pmtech_b200.scene import DeviceScene
    with DeviceScene(64, 0) as dev:
        with pytest.raises(capi.PmbError):
            dev.apply_root_motion([0], [[0, 0, 0, 1]], [[1, 0, 0]])
        bad = np.zeros(1, capi.SKIN_DTYPE)
        bad["num_joints"] = 86
        with pytest.raises(capi.PmbError):
            dev.set_skins(bad, np.zeros((86, 16), np.float32), 86)
        bad["num_joints"], bad["joints_offset"] = 10, 60
        with pytest.raises(capi.PmbError):
            dev.set_skins(bad, np.zeros((10, 16), np.float32), 10)


@pytest.mark.parametrize("world", [1, 3])
def test_peer_push_kernels_match_the_host_statement(gpu, oracle, cams, world):
    """pmb_publish_counts / pmb_push_row (the peer-memory form of the exchange) with the ranks of a box
    simulated on one GPU: every 'rank' owns a table and a list buffer, all reachable from this device,
    as the symmetric-memory mappings are on a real box.  Every rank must end up with the unsharded lists."""
    import torch
    from pmtech_b200 import synth
    from pmtech_b200.scene import DeviceScene
    n = 40_000
    sc = synth.city_scene(n, seed=29)
    views = synth.city_views(cams)
    nv = len(views)
    want = run_oracle(oracle, sc, views)["lists"]
    total = sum(len(l) for l in want)
    parts = synth.partition_by_root(sc["parents"], world)
    stride = (total + nv + 7) & ~3
    cap = stride * world
    rows = torch.zeros(world, stride, dtype=torch.int32, device="cuda")
    tabs = torch.full((world, 16 * 16), -1, dtype=torch.int32, device="cuda")
    outs = torch.zeros(world, cap, dtype=torch.int32, device="cuda")
    metas = torch.zeros(world, 2 * nv + 1, dtype=torch.int32, device="cuda")
    tab_ptrs = [tabs[q].data_ptr() for q in range(world)]
    out_ptrs = [outs[q].data_ptr() for q in range(world)]
    with DeviceScene(n, 0) as dev:
        for r, (first, last) in enumerate(parts):
            upload_scene(dev, synth.slice_scene(sc, first, last))
            dev.frame(last - first, views)
            dev.pack_row(nv, first, rows[r].data_ptr(), stride)
            dev.publish_counts(rows[r].data_ptr(), nv, r, world, tab_ptrs)
            dev.synchronize()
        for r in range(world):  # (on a box: after the barrier)
            dev.push_row(rows[r].data_ptr(), stride, nv, r, world, tabs[r].data_ptr(), out_ptrs, cap, metas[r].data_ptr())
        dev.synchronize()
        for q in range(world):
            m = metas[q].cpu().numpy()
            assert m[2 * nv] == 0
            o = outs[q].cpu().numpy().view(np.uint32)
            for v in range(nv):
                assert np.array_equal(o[m[v]:m[v] + m[nv + v]], want[v]), (q, v)
        # lists that do not fit are reported by every rank and nothing is written
        metas.zero_()
        dev.push_row(rows[0].data_ptr(), stride, nv, 0, world, tabs[0].data_ptr(), out_ptrs, max(total - 1, 0), metas[0].data_ptr())
        dev.synchronize()
        assert metas[0].cpu().numpy()[2 * nv] == 1 or total == 0


@pytest.mark.parametrize("n", [1, 257, 6000])
def test_draw_call_packing(gpu, oracle, cams, n):
    """pmb_update_draw_calls against the oracle's ecs_scene.cpp:1962-1991: world / identity, v1.xy, the
    inverse transpose, and everything the reference leaves alone (v1.zw, v2, records it skips)."""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    sc = synth.mixed_scene(n=n, seed=41)
    rng = np.random.default_rng(5)
    ent = sc["entities"].copy()
    ent[rng.random(n) < 0.05] |= np.uint64(1 << 6)
    ent[rng.random(n) < 0.03] |= np.uint64(1 << 16)
    sc["entities"] = ent
    cb = rng.integers(1, 5000, n).astype(np.uint32)
    cb[rng.random(n) < 0.1] = 0
    cb[rng.random(n) < 0.1] = 0xFFFFFFFF
    dc0 = rng.standard_normal((n, 40)).astype(np.float32)
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], ent, sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc.get("world_matrices"), pos_extent=sc.get("pos_extent"),
                           has_rb_matrix=sc.get("has_rb_matrix"))
    o.update_scene()
    want = o.update_draw_calls(cb, dc0.copy(), 777.25)
    want0 = o.update_draw_calls(cb, np.zeros((n, 40), np.float32), 777.25)  # records the host never touched start zeroed
    with DeviceScene(n, 0) as dev:
        upload_scene(dev, sc)
        dev.upload(capi.CBUFFER, cb)
        dev.update_scene(n)
        dev.update_draw_calls(n, 777.25)  # no PMB_DRAW_CALLS upload yet: the variant that does not read v1 / v2 back
        got0 = dev.download(capi.DRAW_CALLS, 0, n)
        dev.upload(capi.DRAW_CALLS, dc0)
        dev.update_draw_calls(n, 777.25)
        got = dev.download(capi.DRAW_CALLS, 0, n)
    for g, w in ((got0, want0), (got, want)):
        same = (g.view(np.uint32) == w.view(np.uint32)) | (np.isnan(g) & np.isnan(w))  # NaN payloads are not compared
        assert same.all(), np.argwhere(~same)[:5]
    assert np.isfinite(got[:, 24:40]).mean() > 0.9


def test_golden_stage0_fixture_on_device(gpu, cams):
    """the committed animated-rig fixture (oracle output, tests/golden/make_golden.py) replayed on the GPU"""
    import os
    import make_golden_lib as mg
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "rigs_stage0.npz"))
    case = mg.stage0_case(cams)
    got = mg.run_stage0_device(case)
    for k in mg.STAGE0_KEYS:
        a, b = np.ascontiguousarray(got[k]), z[k]
        if a.dtype == np.float32:
            same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
            assert same.all(), (k, np.argwhere(~same)[:4])
        else:
            assert np.array_equal(a, b), k
    for v in range(len(case[1])):
        assert np.array_equal(got["lists"][v], z[f"list{v}"]), v


def test_resize_keeps_the_optional_mirrors(gpu):
    from pmtech_b200 import capi
    from pmtech_b200.scene import DeviceScene
    rng = np.random.default_rng(2)
    n = 1500
    tr = rng.standard_normal((n, 10)).astype(np.float32)
    cb = rng.integers(0, 99, n).astype(np.uint32)
    dc = rng.standard_normal((n, 40)).astype(np.float32)
    with DeviceScene(n, 0) as dev:
        dev.upload(capi.TRANSFORMS, tr)
        dev.upload(capi.CBUFFER, cb)
        dev.upload(capi.DRAW_CALLS, dc)
        dev.synchronize()
        dev.resize(5000)
        assert np.array_equal(dev.download(capi.TRANSFORMS, 0, n), tr)
        assert np.array_equal(dev.download(capi.CBUFFER, 0, n), cb)
        assert np.array_equal(dev.download(capi.DRAW_CALLS, 0, n), dc)
        assert not dev.download(capi.TRANSFORMS, n, 5000 - n).any()  # new memory is zeroed (ecs_scene.cpp:122-152)
        assert not dev.download(capi.DRAW_CALLS, n, 5000 - n).any()


@pytest.mark.parametrize("tail", [True, False])
@pytest.mark.parametrize("hierarchy", ["chains", "wavefront"])
def test_both_forms_of_loop_a_and_both_frame_shapes_give_the_same_bits(gpu, oracle, cams, hierarchy, tail):
    """loop A as re-multiplied chains or as a wavefront over depth levels (PMB_OPT_HIERARCHY), the frame as one launch
    or as separate launches (PMB_OPT_FUSED_TAIL): four ways to schedule the same arithmetic, one set of bits.  (CHAINS
    is honoured up to 16 levels; the depth-40 scene runs the wavefront form either way.)"""
    from pmtech_b200 import synth
    sc = synth.mixed_scene(n=9000, seed=19, max_depth=40)
    views = synth.mixed_views(cams)
    want = run_oracle(oracle, sc, views)
    worst = assert_parity(run_device(sc, views, hierarchy=hierarchy, tail=tail), want, f"mixed {hierarchy} tail={tail}")
    assert max(worst.values()) == 0, worst
    assert_parity(run_device(sc, views, mode="sphere", hierarchy=hierarchy, tail=tail), run_oracle(oracle, sc, views, mode="sphere"),
                  f"mixed sphere {hierarchy} tail={tail}")
    deep = synth.deep_scene(n=50_000, max_depth=16, seed=2, block=4096)
    worst = assert_parity(run_device(deep, None, hierarchy=hierarchy, tail=tail), run_oracle(oracle, deep, None),
                          f"deep {hierarchy} tail={tail}")
    assert max(worst.values()) == 0, worst
    shallow = synth.city_scene(70_001, seed=8)
    cv = synth.city_views(cams)
    worst = assert_parity(run_device(shallow, cv, hierarchy=hierarchy, tail=tail), run_oracle(oracle, shallow, cv),
                          f"city {hierarchy} tail={tail}")
    assert max(worst.values()) == 0, worst
    rigs = synth.rig_scene(300, seed=9)
    worst = assert_parity(run_device(rigs, views, mode="sphere", hierarchy=hierarchy, tail=tail),
                          run_oracle(oracle, rigs, views, mode="sphere"), f"rigs {hierarchy} tail={tail}")
    assert max(worst.values()) == 0, worst


def test_a_chain_as_long_as_the_scene_runs_as_a_wavefront(gpu, oracle):
    """every entity the child of the one before it: each tile waits for the tile before it, level by level"""
    from pmtech_b200 import synth
    n = 3000
    rng = np.random.default_rng(5)
    sc = synth.mixed_scene(n=n, seed=5)
    sc["parents"] = np.maximum(np.arange(n, dtype=np.int64) - 1, 0).astype(np.uint32)
    q = synth.random_unit_quats(rng, n)
    sc["local_matrices"] = synth.trs_matrices(rng.uniform(-1, 1, (n, 3)).astype(np.float32), q, np.ones((n, 3), np.float32))
    want = run_oracle(oracle, sc, None)
    for tail in (True, False):
        worst = assert_parity(run_device(sc, None, tail=tail), want, f"chain tail={tail}")
        assert max(worst.values()) == 0, worst


def test_a_chain_too_long_for_ancestor_lists_walks_the_parent_links(gpu, oracle):
    """40,000 entities in one chain: 625 boundary children x depth 40,000 is past the budget of the flat ancestor
    lists (their size grows with the square of a chain), so loop C across tiles walks parent links instead — same
    boxes, also when the one-launch frame is asked for"""
    from pmtech_b200 import synth
    n = 40_000
    rng = np.random.default_rng(6)
    sc = synth.mixed_scene(n=n, seed=6)
    sc["parents"] = np.maximum(np.arange(n, dtype=np.int64) - 1, 0).astype(np.uint32)
    q = synth.random_unit_quats(rng, n)
    sc["local_matrices"] = synth.trs_matrices(rng.uniform(-1, 1, (n, 3)).astype(np.float32), q, np.ones((n, 3), np.float32))
    want = run_oracle(oracle, sc, None)
    for tail in (True, False):
        worst = assert_parity(run_device(sc, None, tail=tail), want, f"long chain tail={tail}")
        assert max(worst.values()) == 0, worst


def test_the_tile_trace_follows_a_resize(gpu, oracle):
    """PMB_OPT_TRACE set on a small handle that then grows: the per-tile timeline is re-allocated with the mirrors"""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    sc = synth.deep_scene(n=150_000, max_depth=12, seed=9)
    n = sc["num_entities"]
    want = run_oracle(oracle, sc, None)
    with DeviceScene(2048) as dev:
        dev.set_option(capi.OPT_TRACE, 1)
        dev.resize(n)
        upload_scene(dev, sc)
        dev.update_scene(n)
        tiles = (n + 63) // 64
        tr = dev.download_trace(tiles)
        assert (tr[:, 1] >= tr[:, 0]).all() and (tr[:, 0] > 0).all()
        assert ulp_diff(dev.download(capi.WORLD_MATRICES, 0, n), want["world"]) == 0


def test_reuploading_unchanged_parents_keeps_the_hierarchy_index(gpu, oracle, cams):
    """a binding without dirty tracking re-sends parents every frame: same values = no rebuild, new values = rebuild"""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene, download_outputs
    sc = synth.mixed_scene(n=5000, seed=23, max_depth=12)
    views = synth.mixed_views(cams)
    n = sc["num_entities"]
    with DeviceScene(n) as dev:
        upload_scene(dev, sc)
        dev.frame(n, views)
        aux0 = dev.kernel_stats()["aux"][1]
        upload_scene(dev, sc)  # flags were cleared by the first frame: the second upload restores them too
        dev.frame(n, views)
        aux1 = dev.kernel_stats()["aux"][1]
        assert aux1 - aux0 <= 3, "unchanged parents must not rebuild the hierarchy index"
        assert_parity(download_outputs(dev, n, len(views)), run_oracle(oracle, sc, views), "re-upload")
        sc2 = dict(sc)
        sc2["parents"] = sc["parents"].copy()
        kids = np.flatnonzero(sc["parents"] != np.arange(n))
        sc2["parents"][kids[::3]] = np.arange(n, dtype=np.uint32)[kids[::3]]  # a third of the children become roots
        upload_scene(dev, sc2)
        dev.frame(n, views)
        assert_parity(download_outputs(dev, n, len(views)), run_oracle(oracle, sc2, views), "changed parents")


@pytest.mark.gpu
@pytest.mark.parametrize("first,count", [(0, None), (7, 1001), (66, 1), (1, 0)])
def test_marking_a_transform_range_dirty_equals_uploading_the_flags(first, count):
    """pmb_mark_transforms_dirty == the host OR-ing e_cmp::transform into entities[first:first+count] and uploading them"""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    from helpers import upload_scene
    sc = synth.rig_scene(20, seed=3)
    n = sc["num_entities"]
    count = n - first if count is None else count
    views = synth.mixed_views(__import__("pmtech_b200.ecs", fromlist=["Camera"]).Camera)
    outs = []
    for on_device in (False, True):
        ent = sc["entities"].copy()
        if not on_device:
            ent[first:first + count] |= np.uint64(1 << 12)
        with DeviceScene(n, 0) as dev:
            upload_scene(dev, dict(sc, entities=ent))
            if on_device:
                dev.mark_transforms_dirty(first, count)
                assert np.array_equal(dev.download(capi.ENTITIES, 0, n) != sc["entities"],
                                      (np.arange(n) >= first) & (np.arange(n) < first + count) & ((sc["entities"] & np.uint64(1 << 12)) == 0))
            dev.frame(n, views, "sphere")
            outs.append([dev.download(a, 0, n) for a in (capi.LOCAL_MATRICES, capi.WORLD_MATRICES, capi.ENTITIES, capi.TRANSFORMS,
                                                          capi.BOUNDING_VOLUMES)] + dev.lists(len(views)))
    for a, b in zip(*outs):
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8))
    with DeviceScene(16, 0) as dev:
        with pytest.raises(capi.PmbError):
            dev.mark_transforms_dirty(10, 100)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 31, 33, 1025, 262144, 262145])
def test_draw_lists_on_both_sides_of_the_small_scene_compaction(oracle, n):
    """scenes up to 256 Ki entities compact their visibility bitmaps in one-warp chunks of 4 Ki entities, larger ones
    in 32 Ki chunks: same ascending lists on either side of the switch"""
    from pmtech_b200 import synth
    sc = synth.mixed_scene(n=n, seed=n % 97)
    views = synth.mixed_views(oracle.OracleCamera)
    want = run_oracle(oracle, sc, views, mode="aabb")
    for fused in (True, False):
        got = run_device(sc, views, mode="aabb", fused=fused)
        assert len(got["lists"]) == len(views)
        for v in range(len(views)):
            assert np.array_equal(got["lists"][v], want["lists"][v]), (n, fused, v)
    assert n < 64 or sum(len(l) for l in want["lists"]) > 0


@pytest.mark.gpu
def test_a_refused_exchange_launch_leaves_the_previous_frame_to_be_completed(oracle, cams):
    """pmb_exchange_launch checks what it can refuse (here: the last cull built fewer views than the exchange carries)
    before it touches its state: the frame launched before stays pending and is completed by the next launch; the
    lists after the refusal are still the oracle's, on both buffer sets."""
    import ctypes as C
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    L = capi.load()
    sc = synth.city_scene(50_000, seed=5)
    views = synth.city_views(cams)
    n, nv = sc["num_entities"], len(views)
    want = run_oracle(oracle, sc, synth.city_views(oracle.OracleCamera))["lists"]
    with DeviceScene(n) as dev:
        x = C.c_void_p()
        capi._check(L.pmb_exchange_create(C.byref(x), dev._h, 0, 1, nv, 0, n * nv + nv + 64, n * nv + 64, capi.XCHG_ROOT, 0))
        try:
            arr = (C.c_void_p * 1)(x)
            capi._check(L.pmb_exchange_connect_local(arr, 1))
            for step in range(4):
                upload_scene(dev, sc)
                dev.frame(n, views)
                capi._check(L.pmb_exchange_launch(x))
                if step == 1:
                    dev.cull_views(n, views[:3])  # the scene now holds 3 lists: the exchange wants 11
                    assert L.pmb_exchange_launch(x) == -5 and "views" in L.pmb_last_error().decode()  # PMB_ERR_STATE
                    assert L.pmb_exchange_launch(x) == -5  # and again: still nothing touched
                    continue
                res = capi.ExchangeResult()
                if step == 2:
                    continue  # two launches back to back: frame 1's phases ride behind frame 2's publication
                capi._check(L.pmb_exchange_finish(x, C.byref(res)))
                assert res.overflow == 0 and res.world == 1 and res.num_views == nv
                for v in range(nv):
                    buf = np.empty(len(want[v]) + 1, np.uint32)
                    got = L.pmb_exchange_get_list(x, v, capi._ptr(buf), len(buf))
                    assert got == len(want[v]) and np.array_equal(buf[:got], want[v]), (step, v)
        finally:
            L.pmb_exchange_destroy(x)


@pytest.mark.gpu
def test_group_exchange_with_polling_kernels(oracle, cams):
    """the fallback for drivers without stream memory operations: the same protocol, the waits as one-warp kernels"""
    from pmtech_b200 import capi
    L = capi.load()
    L.pmb_exchange_use_polling_kernels(1)
    try:
        test_group_exchange_gathers_the_unsharded_lists(oracle, cams, 3, "root")
    finally:
        L.pmb_exchange_use_polling_kernels(0)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["root", "all", "counts"])
@pytest.mark.parametrize("world", [1, 2, 3])
def test_group_exchange_gathers_the_unsharded_lists(oracle, cams, world, mode):
    """pmb_group (one process, `world` root-aligned shards; on a one-GPU box they share the GPU and run the same
    device-flag barrier and peer-store kernels): counts table, renderable extents over all shards and the gathered
    global lists equal the oracle's for the unsharded scene, on both buffer sets."""
    import torch
    from pmtech_b200 import capi, shard, synth
    n = 120_000
    sc = synth.city_scene(n, seed=31)
    views = synth.city_views(cams)
    nv = len(views)
    want = run_oracle(oracle, sc, synth.city_views(oracle.OracleCamera))
    parts = synth.partition_by_root(sc["parents"], world)
    ndev = capi.device_count()
    with shard.DeviceGroup([r % ndev for r in range(world)], [b - a for a, b in parts]) as g:
        for r, (a, b) in enumerate(parts):
            upload_scene(g.scenes[r], synth.slice_scene(sc, a, b))
        total_want = sum(len(l) for l in want["lists"])
        g.exchange_setup(nv, [a for a, _ in parts], total_want + nv + 64, total_want + 64, mode=mode, root=world - 1)
        for rep in range(3):  # two frames back to back (the second one's publication overtakes the first one's push), then one more
            g.frame([b - a for a, b in parts], views)
            if rep:
                res = g.finish(rank=world - 1)
        assert not res["overflow"] and res["epoch"] == 3
        assert np.array_equal(res["view_totals"], [len(l) for l in want["lists"]])
        assert np.array_equal(res["counts"].sum(axis=0), res["view_totals"])
        assert np.array_equal(res["extents"].view(np.uint32), np.asarray(want["extents"], np.float32).view(np.uint32))
        if mode == "counts":
            assert res["lists_ptr"] == 0
        else:
            x = g._L.pmb_group_exchange(g._g, world - 1)
            for v in range(nv):
                buf = np.zeros(max(len(want["lists"][v]), 1), np.uint32)
                c = g._L.pmb_exchange_get_list(x, v, capi._ptr(buf), buf.shape[0])
                assert c == len(want["lists"][v]), (v, c, g._L.pmb_last_error())
                assert np.array_equal(buf[:c], want["lists"][v]), v
            other = g.finish(rank=0)
            assert (other["lists_ptr"] != 0) == (mode == "all" or world == 1)
        # a row that cannot hold the lists is reported, nothing is pushed
        g.exchange_setup(nv, [a for a, _ in parts], nv + 8, total_want + 64, mode=mode)
        g.frame([b - a for a, b in parts], views)
        assert g.finish(rank=0)["overflow"] == (mode != "counts")  # no lists travel in "counts": no row to overflow


@pytest.mark.gpu
def test_instance_blocks_of_a_master_sub_instance_scene(gpu, oracle, cams):
    """ecs_scene.cpp:1962-2011 for an instanced scene (examples/code/instancing): the masters' records are packed, the
    sub-instances' records keep their inverse transpose but get world / v1.xy, and each master's instance buffer is the
    run of records behind it — on the device, bit for bit the oracle's draw_call_data[n + 1 ..]."""
    from pmtech_b200 import capi, synth
    from pmtech_b200.scene import DeviceScene
    rng = np.random.default_rng(21)
    sc = synth.mixed_scene(n=3000, seed=21)
    n = sc["num_entities"]
    ent = sc["entities"].copy()
    mi = np.zeros(n, capi.MASTER_INSTANCE_DTYPE)
    masters = {100: 64, 400: 1, 1000: 300, 2990: 9}
    for m, k in masters.items():
        ent[m] = np.uint64(0x13 | (1 << 15))
        ent[m + 1:m + 1 + k] = np.uint64(0x13 | (1 << 14))
        mi[m] = (k, 500 + m, 160)
    sc = dict(sc, entities=ent)
    cb = rng.integers(1, 1000, n).astype(np.uint32)
    dc0 = rng.standard_normal((n, 40)).astype(np.float32)
    o = oracle.OracleScene(sc["parents"], sc["local_matrices"], ent, sc["state_flags"], sc["bounding_volumes"],
                           world_matrices=sc["world_matrices"], pos_extent=sc["pos_extent"], has_rb_matrix=sc["has_rb_matrix"])
    o.update_scene()
    want = o.update_draw_calls(cb, dc0.copy(), 77.0)
    L = capi.load()
    blocks = np.zeros(8, capi.INSTANCE_BLOCK_DTYPE)
    cnt = L.pmb_instance_blocks(capi._ptr(ent), capi._ptr(mi), n, capi._ptr(blocks), 8)
    assert cnt == 4 and [int(b["master"]) for b in blocks[:cnt]] == sorted(masters)
    with DeviceScene(n) as dev:
        upload_scene(dev, sc)
        dev.upload(capi.CBUFFER, cb)
        dev.upload(capi.DRAW_CALLS, dc0)
        dev.update_scene(n)
        dev.update_draw_calls(n, 77.0)
        assert L.pmb_device_draw_calls(dev._h)
        for b in blocks[:cnt]:
            m, k = int(b["master"]), masters[int(b["master"])]
            assert int(b["bytes"]) == k * 160 and int(b["byte_offset"]) == (m + 1) * 160
            got = np.zeros((k, 40), np.float32)
            one = np.array([b], capi.INSTANCE_BLOCK_DTYPE)
            capi._check(L.pmb_download_instance_block(dev._h, capi._ptr(one), capi._ptr(got)))
            assert np.array_equal(got.view(np.uint32), want[m + 1:m + 1 + k].view(np.uint32)), m
            # a sub-instance record: world + v1.xy rewritten, the inverse transpose left as it was (:1981-1982)
            assert np.array_equal(got[:, 24:40].view(np.uint32), dc0[m + 1:m + 1 + k, 24:40].view(np.uint32))
            assert np.array_equal(got[:, 16], np.arange(m + 1, m + 1 + k, dtype=np.float32))
        bad = np.array([(0, 0, (n + 5000) * 160, 160)], capi.INSTANCE_BLOCK_DTYPE)
        assert L.pmb_download_instance_block(dev._h, capi._ptr(bad), capi._ptr(np.zeros(40, np.float32))) == -1
