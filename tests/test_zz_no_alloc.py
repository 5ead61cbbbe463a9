"""pmb_cull_views_static — the cull + draw-list build for hosts that may not allocate GPU memory.

Who needs it: an engine that gave its whole GPU memory budget to its own pools at start-up (a later cudaMalloc fails or
eats what the renderer had counted on), code in the middle of a frame (cudaMalloc / cudaFree synchronise the device and
take an unbounded time), a plugin let into somebody else's CUDA context on the condition that it allocates nothing.
They cannot create a pmb_scene (it cudaMalloc's mirrors, bitmaps, counters and lists), so this entry point takes
every buffer from the caller and keeps its scratch in a static device array inside the library (about 2 MiB per GPU,
up to PMB_STATIC_MAX_ENTITIES = 1 Mi entities and PMB_MAX_VIEWS = 16 views per call).

What is checked, through the C ABI: its lists and counts equal those of the allocating entry point, pmb_cull_views,
on the same device arrays — bit for bit, the tolerance every draw-list test here uses — on two calls in a row that
share the static scratch (the second smaller, in another cull mode, with fewer views: what the first call left in the
scratch must not show), and once more with the first call's arguments.  Oversized requests are refused by name.

The file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should not hide the parity
suite.  It holds no state of its own, so it passes alone in a fresh process and behind the other tests alike.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import upload_scene

pytestmark = pytest.mark.gpu

STATIC_MAX_ENTITIES = 1 << 20  # PMB_STATIC_MAX_ENTITIES
_MODES = {"aabb": 0, "sphere": 1, "filter": 2}


def _static_cull(lib, dev, n, views, mode, lists_t, counts_t, stride):
    """one call of the no-allocation entry point on the scene's own device arrays; outputs are the caller's"""
    from pmtech_b200 import capi
    fr = np.ascontiguousarray(views, np.float32).reshape(-1, capi.FRUSTUM_FLOATS)
    rc = lib.pmb_cull_views_static(C.c_void_p(dev.device_ptr(capi.ENTITIES)), C.c_void_p(dev.device_ptr(capi.STATE_FLAGS)),
                                   C.c_void_p(dev.device_ptr(capi.POS_EXTENT)), n, C.c_void_p(fr.ctypes.data), fr.shape[0],
                                   _MODES[mode], C.c_void_p(lists_t.data_ptr()), stride, C.c_void_p(counts_t.data_ptr()), None)
    assert rc == 0, lib.pmb_last_error().decode()
    return fr.shape[0]


def _compare(dev, n, views, mode, lists_t, counts_t, stride, what):
    """the existing entry point (pmb_cull_views -> pmb_get_counts / pmb_get_list) against the caller's buffers"""
    import torch
    nv = dev.cull_views(n, views, mode)
    want = dev.lists(nv)
    torch.cuda.synchronize()
    counts = counts_t.cpu().numpy().view(np.uint32)
    lists = lists_t.cpu().numpy().view(np.uint32).reshape(-1, stride)
    for v in range(nv):
        assert counts[v] == len(want[v]), f"{what} view {v}: {counts[v]} entries, pmb_cull_views {len(want[v])}"
        assert np.array_equal(lists[v, :counts[v]], want[v]), f"{what} view {v}: draw list differs"
    return [len(w) for w in want]


def test_static_scratch_cull_equals_the_allocating_cull_on_consecutive_calls(gpu):
    import torch
    from pmtech_b200 import synth
    from pmtech_b200.ecs import Camera
    from pmtech_b200.scene import DeviceScene
    n = STATIC_MAX_ENTITIES  # the largest scene the built-in scratch takes
    sc = synth.city_scene(n)
    views = synth.city_views(Camera)  # 11 views: camera, 4 cascades, 6 cube faces
    dev_t = torch.device("cuda:0")
    with DeviceScene(n) as dev:
        upload_scene(dev, sc)
        dev.update_scene(n)  # loop B leaves pos_extent on the device: the cull's input
        dev.synchronize()
        stride = n
        # the caller's output buffers (allocated here, by the test, with torch: the entry point itself allocates nothing)
        lists_t = torch.full((len(views) * stride,), -1, dtype=torch.int32, device=dev_t)
        counts_t = torch.full((16,), -1, dtype=torch.int32, device=dev_t)
        torch.cuda.synchronize()

        # call 1: full size, 11 views, box test
        _static_cull(gpu, dev, n, views, "aabb", lists_t, counts_t, stride)
        sizes = _compare(dev, n, views, "aabb", lists_t, counts_t, stride, "call 1")
        assert 0 < sum(sizes) < n * len(views)

        # call 2, straight after: fewer entities (not a multiple of 32), fewer views, sphere test — the scratch still
        # holds call 1's bitmaps and chunk offsets beyond what this call writes
        n2, views2 = 300_001, views[[4, 0, 7]]
        lists_t.fill_(-1)
        _static_cull(gpu, dev, n2, views2, "sphere", lists_t, counts_t, stride)
        sizes2 = _compare(dev, n2, views2, "sphere", lists_t, counts_t, stride, "call 2")
        assert 0 < sum(sizes2) < n2 * 3

        # and call 1 again behind the smaller one
        _static_cull(gpu, dev, n, views, "aabb", lists_t, counts_t, stride)
        assert _compare(dev, n, views, "aabb", lists_t, counts_t, stride, "call 3") == sizes

        # filter only: no planes, pos_extent not needed
        from pmtech_b200 import capi
        fr = np.ascontiguousarray(views[:1], np.float32)
        rc = gpu.pmb_cull_views_static(C.c_void_p(dev.device_ptr(capi.ENTITIES)), C.c_void_p(dev.device_ptr(capi.STATE_FLAGS)), None,
                                       5000, C.c_void_p(fr.ctypes.data), 1, _MODES["filter"], C.c_void_p(lists_t.data_ptr()), stride,
                                       C.c_void_p(counts_t.data_ptr()), None)
        assert rc == 0, gpu.pmb_last_error().decode()
        _compare(dev, 5000, views[:1], "filter", lists_t, counts_t, stride, "filter only")


def test_requests_beyond_the_built_in_scratch_are_refused(gpu):
    fr = np.zeros((17, 60), np.float32)
    one = C.c_void_p(256)  # never dereferenced: the sizes are checked first
    rc = gpu.pmb_cull_views_static(one, one, one, STATIC_MAX_ENTITIES + 1, C.c_void_p(fr.ctypes.data), 1, 0, one, STATIC_MAX_ENTITIES + 1,
                                   one, None)
    assert rc == -1 and "PMB_STATIC_MAX_ENTITIES" in gpu.pmb_last_error().decode()
    rc = gpu.pmb_cull_views_static(one, one, one, 1000, C.c_void_p(fr.ctypes.data), 17, 0, one, 1000, one, None)
    assert rc == -1 and "17 views" in gpu.pmb_last_error().decode()
    rc = gpu.pmb_cull_views_static(one, one, one, 1000, C.c_void_p(fr.ctypes.data), 2, 0, one, 999, one, None)
    assert rc == -1 and "list_stride" in gpu.pmb_last_error().decode()
