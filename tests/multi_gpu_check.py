#!/usr/bin/env python
"""Multi-GPU check of the list exchange, one process per GPU (launched by tests/test_multi_gpu.py under pytest -m gpu
when the box has two or more GPUs, or by hand):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_check.py

Every rank culls its root-aligned shard of ONE scene (synth.partition_by_root), exchanges with the library's own
peer-memory exchange (lists gathered on rank 0, replicated, counts only) and with the NCCL fallback, and compares every
result with the oracle's lists / extents of the unsharded scene."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from helpers import run_oracle, upload_scene
    from oracle import oracle as orc
    from pmtech_b200 import shard, synth
    from pmtech_b200.ecs import Camera
    from pmtech_b200.scene import DeviceScene
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n = int(os.environ.get("PMB_CHECK_ENTITIES", 300_000))
    sc = synth.city_scene(n, seed=31)
    views = synth.city_views(Camera)
    nv = len(views)
    ref = run_oracle(orc, sc, synth.city_views(orc.OracleCamera))
    want, want_ext = ref["lists"], np.asarray(ref["extents"], np.float32)
    first, last = synth.partition_by_root(sc["parents"], world)[rank]
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ok = True
    with DeviceScene(last - first, local_rank) as dev:
        dev.set_stream(stream.cuda_stream)
        upload_scene(dev, synth.slice_scene(sc, first, last))
        dev.frame(last - first, views)
        kinds = [("peer-root", lambda: shard.Exchange(dev, nv, first, last - first, scene_stream=stream, mode="root")),
                 ("peer-all", lambda: shard.Exchange(dev, nv, first, last - first, scene_stream=stream, mode="all")),
                 ("peer-counts", lambda: shard.Exchange(dev, nv, first, last - first, scene_stream=stream, mode="counts")),
                 ("nccl", lambda: shard.ListGather(dev, nv, first, last - first, scene_stream=stream))]
        for kind, make in kinds:
            g = make()
            for rep in range(3):  # both buffer sets and a reuse
                dev.frame(last - first, views)
                g.launch()
            total = g.finish()
            good = total == sum(len(l) for l in want)
            if kind.startswith("peer"):
                r = g.result
                good &= bool(np.array_equal(r["extents"].view(np.uint32), want_ext.view(np.uint32)))
                good &= bool(np.array_equal(r["counts"].sum(axis=0), [len(l) for l in want]))
            holds = kind in ("peer-all", "nccl") or (kind == "peer-root" and rank == 0)
            if holds:
                for v in range(nv):
                    got = g.global_list(v).cpu().numpy().view(np.uint32)
                    if not np.array_equal(got, want[v]):
                        good = False
                        print(f"rank {rank} {kind}: view {v} differs ({got.shape[0]} vs {want[v].shape[0]} entries)", flush=True)
            ok &= good
            print(f"rank {rank} {kind}: {total} entries, row_stride {g.row_stride}, overflows {g.overflows}, "
                  f"holds lists {holds}: {'ok' if good else 'MISMATCH'}", flush=True)
            torch.cuda.synchronize()
            dist.barrier()
            if hasattr(g, "close"):
                g.close()
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if rank == 0:
        print("EXCHANGE CHECK", "PASSED" if int(t.item()) else "FAILED", flush=True)
    sys.exit(0 if int(t.item()) else 1)


if __name__ == "__main__":
    main()
