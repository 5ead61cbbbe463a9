"""debug variants: argv[1] = variant"""
import os, sys, time, ctypes as C
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from pmtech_b200 import capi, shard, synth
from pmtech_b200.ecs import Camera
variant = sys.argv[1]
world = 2
n = 60000
sc = synth.city_scene(n, seed=31)
views = synth.city_views(Camera)
parts = synth.partition_by_root(sc["parents"], world)
L = capi.load()
L.pmb_group_exchange.restype = C.c_void_p
with shard.DeviceGroup([0] * world, [b - a for a, b in parts]) as g:
    for r, (a, b) in enumerate(parts):
        sub = synth.slice_scene(sc, a, b)
        d = g.scenes[r]
        d.set_option(capi.OPT_FUSED_TAIL, 0)
        for arr, k in ((capi.PARENTS, "parents"), (capi.LOCAL_MATRICES, "local_matrices"), (capi.ENTITIES, "entities"),
                       (capi.STATE_FLAGS, "state_flags"), (capi.BOUNDING_VOLUMES, "bounding_volumes")):
            d.upload(arr, sub[k])
        d.synchronize()
    mode = "counts" if variant == "counts" else "root"
    g.exchange_setup(len(views), [a for a, _ in parts], 11 * n, 11 * n, mode=mode, root=0)
    xs = [C.c_void_p(L.pmb_group_exchange(g._g, r)) for r in range(world)]
    for r, (a, b) in enumerate(parts):
        g.scenes[r].frame(b - a, views)
    if variant != "nosync":
        for r in range(world):
            g.scenes[r].synchronize()
    print("frames done", flush=True)
    order = [1, 0] if variant == "reverse" else [0, 1]
    for r in order:
        rc = L.pmb_exchange_launch(xs[r])
        print("launch", r, rc, flush=True)
        if variant == "sleep":
            time.sleep(1.0)
    if variant == "nowaits":
        time.sleep(1.0)
        L.pmb_exchange_debug_header.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        for r in range(world):
            hdr = np.zeros(1248, np.uint32)
            rc = L.pmb_exchange_debug_header(xs[r], capi._ptr(hdr), 1248)
            print("rank", r, "rc", rc, "READY", hdr[1024:1024 + 32].reshape(2, 16)[:, :world].tolist(), "DONE", hdr[1056:1056 + 32].reshape(2, 16)[:, :world].tolist(),
                  "tab counts", hdr[0:1024].reshape(2, 16, 32)[:, :world, :3].tolist(), "ctr", hdr[1216:1220].tolist(), flush=True)
    res = g.finish(0)
    print("OK", variant, res["view_totals"].tolist(), flush=True)
