"""debug: two shards on one GPU through pmb_group; argv: world tail(0/1/2) frames"""
import os, sys
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from pmtech_b200 import capi, shard, synth
from pmtech_b200.ecs import Camera
world, tail, frames = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
n = 60000
sc = synth.city_scene(n, seed=31)
views = synth.city_views(Camera)
parts = synth.partition_by_root(sc["parents"], world)
with shard.DeviceGroup([0] * world, [b - a for a, b in parts]) as g:
    for r, (a, b) in enumerate(parts):
        sub = synth.slice_scene(sc, a, b)
        d = g.scenes[r]
        d.set_option(capi.OPT_FUSED_TAIL, tail)
        for arr, k in ((capi.PARENTS, "parents"), (capi.LOCAL_MATRICES, "local_matrices"), (capi.ENTITIES, "entities"),
                       (capi.STATE_FLAGS, "state_flags"), (capi.BOUNDING_VOLUMES, "bounding_volumes")):
            d.upload(arr, sub[k])
        d.synchronize()
    g.exchange_setup(len(views), [a for a, _ in parts], 11 * n, 11 * n, mode="root", root=0)
    for f in range(frames):
        g.frame([b - a for a, b in parts], views)
        print("enqueued", f, flush=True)
        import time, ctypes as C
        time.sleep(0.1)
        L = g._L
        L.pmb_exchange_debug_header.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        for r in range(world):
            hdr = np.zeros(1248, np.uint32)
            L.pmb_group_exchange.restype = C.c_void_p
            x = L.pmb_group_exchange(g._g, r)
            rc = L.pmb_exchange_debug_header(C.c_void_p(x), capi._ptr(hdr), 1248)
            print("rank", r, "rc", rc, "READY", hdr[1024:1024 + 32].reshape(2, 16)[:, :world].tolist(), "DONE", hdr[1056:1056 + 32].reshape(2, 16)[:, :world].tolist(),
                  "tab counts", hdr[0:1024].reshape(2, 16, 32)[:, :world, :3].tolist(), "ctr", hdr[1216:1220].tolist(), flush=True)
        res = g.finish(0)
        print("frame", f, res["view_totals"].tolist(), res["device_ms"], flush=True)
print("OK")
