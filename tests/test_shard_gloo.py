"""CPU suite, part 3: the N > 1 path over gloo with world_size 2 (and 3).

Each rank owns a root-aligned shard (pmtech_b200.synth.partition_by_root), produces its per-view
lists, and the exchange of pmtech_b200.shard (counts all_gather -> packed all_gather -> segment
table) must rebuild, on every rank, exactly the lists of the unsharded scene.  On the GPU box the
per-shard lists come from the CUDA path and the pack / unpack steps are the segment-copy kernel;
here the lists come from the oracle and the two copies are numpy, so what is under test is the
host logic of the exchange: offsets, padding, rank order, index rebasing, extents reduction."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _apply_segments(src, seg, total):
    dst = np.zeros(total, src.dtype)
    for s in seg:
        c = int(s["count"])
        dst[int(s["dst_offset"]):int(s["dst_offset"]) + c] = src[int(s["src_offset"]):int(s["src_offset"]) + c] + int(s["add"])
    return dst


def _worker(rank, world, port, n, seed, q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch
        import torch.distributed as dist
        from helpers import run_oracle
        from oracle import oracle as orc
        from pmtech_b200 import shard, synth
        dist.init_process_group("gloo", rank=rank, world_size=world)
        sc = synth.city_scene(n, seed=seed)
        views = synth.city_views(orc.OracleCamera)
        nv = len(views)
        first, last = synth.partition_by_root(sc["parents"], world)[rank]
        mine = run_oracle(orc, synth.slice_scene(sc, first, last), views)
        counts = torch.tensor([len(l) for l in mine["lists"]], dtype=torch.int32)
        counts_all = shard.exchange_counts(counts).numpy()
        assert counts_all.shape == (world, nv)
        row_stride = (int(counts_all.sum(axis=1).max()) + 3) & ~3
        # pack: the V lists back to back, local -> global entity index (pmb_pack_lists on the GPU)
        packed = np.zeros(row_stride, np.int32)
        o = 0
        for l in mine["lists"]:
            packed[o:o + len(l)] = l.astype(np.int64) + first
            o += len(l)
        gathered = shard.exchange_packed(torch.from_numpy(packed), row_stride).numpy()
        seg, view_off, view_tot = shard.unpack_segments(counts_all, row_stride)
        out = _apply_segments(gathered, seg, int(view_tot.sum()))
        # the sync-free form: fixed-size rows {counts, lists}, one all_gather, offsets from the headers
        stride2 = int(counts_all.sum(axis=1).max()) + nv + 5
        row = shard.pack_row_np(mine["lists"], first, stride2)
        rows = shard.exchange_packed(torch.from_numpy(row.view(np.int32)), stride2).numpy().view(np.uint32).reshape(world, stride2)
        lists2, overflow = shard.unpack_rows_np(rows, nv)
        assert not overflow
        small = shard.pack_row_np(mine["lists"], first, nv + 1)  # a row that cannot hold its lists is reported, not trusted
        rows_small = shard.exchange_packed(torch.from_numpy(small.view(np.int32)), nv + 1).numpy().view(np.uint32).reshape(world, nv + 1)
        assert shard.unpack_rows_np(rows_small, nv)[1]
        ext = torch.from_numpy(np.asarray(mine["extents"], np.float32).copy())
        mn, mx = shard.reduce_extents(ext)
        res = dict(rank=rank, lists=[out[int(view_off[v]):int(view_off[v]) + int(view_tot[v])].astype(np.uint32)
                                     for v in range(nv)],
                   lists2=lists2, ext=np.concatenate([mn.numpy(), mx.numpy()]), shard=(first, last))
        dist.barrier()
        dist.destroy_process_group()
        q.put(res)
    except Exception as e:  # pragma: no cover
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc() + str(e)))


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_lists_equal_the_unsharded_scene(oracle, world):
    import torch.multiprocessing as mp
    from helpers import run_oracle
    from pmtech_b200 import synth
    n, seed = 60_000, 17
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for r in results:
        assert "error" not in r, r.get("error")
    sc = synth.city_scene(n, seed=seed)
    views = synth.city_views(oracle.OracleCamera)
    want = run_oracle(oracle, sc, views)
    assert sum(len(l) for l in want["lists"]) > 0
    for r in results:
        for v in range(len(views)):
            assert np.array_equal(r["lists"][v], want["lists"][v]), (r["rank"], v)
            assert np.array_equal(r["lists2"][v], want["lists"][v]), (r["rank"], v, "row protocol")
        assert np.array_equal(r["ext"].view(np.uint32), np.asarray(want["extents"], np.float32).view(np.uint32))
    shards = sorted(r["shard"] for r in results)
    assert shards[0][0] == 0 and shards[-1][1] == n


def test_unpack_segments_layout():
    from pmtech_b200 import shard
    counts = np.array([[2, 0, 3], [1, 4, 0]])
    seg, off, tot = shard.unpack_segments(counts, row_stride=8)
    assert tot.tolist() == [3, 4, 3] and off.tolist() == [0, 3, 7]
    table = {(int(s["src_offset"]), int(s["dst_offset"]), int(s["count"])) for s in seg}
    # rank 0 row starts at 0: view0 @0 (2), view1 @2 (0), view2 @2 (3); rank 1 row starts at 8
    assert (0, 0, 2) in table and (2, 7, 3) in table and (8, 2, 1) in table and (9, 3, 4) in table


def test_row_protocol_layout_and_overflow():
    from pmtech_b200 import shard
    a = shard.pack_row_np([np.array([1, 5]), np.array([], np.uint32), np.array([2, 3, 9])], 100, 12)
    assert a.tolist() == [2, 0, 3, 101, 105, 102, 103, 109, 0, 0, 0, 0]
    b = shard.pack_row_np([np.array([7]), np.array([0, 1, 2, 3]), np.array([], np.uint32)], 200, 12)
    lists, overflow = shard.unpack_rows_np(np.stack([a, b]), 3)
    assert not overflow
    assert [l.tolist() for l in lists] == [[101, 105, 207], [200, 201, 202, 203], [102, 103, 109]]
    c = shard.pack_row_np([np.arange(20)], 0, 8)  # truncated body, true count in the header
    assert c[0] == 20 and c[1:].tolist() == list(range(7))
    assert shard.unpack_rows_np(np.stack([c, c]), 1) == (None, True)


# ---- bench.py's once-per-run check of the exchange, on gloo with stand-ins for the scene and the exchange ----------
def _check_worker(rank, world, port, q, corrupt):
    try:
        sys.path.insert(0, ROOT)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch
        import torch.distributed as dist
        import bench
        dist.init_process_group("gloo", rank=rank, world_size=world)
        n, nv = 1000, 3
        rng = np.random.default_rng(5)
        lists = [[np.sort(rng.choice(n, size=int(rng.integers(0, 200)), replace=False)).astype(np.int32) for _ in range(nv)]
                 for _ in range(world)]
        ext = [np.array([-1 - r, -2, -3, 4, 5 + r, 6], np.float32) for r in range(world)]
        counts_tab = np.array([[len(l) for l in row] for row in lists], np.int64)

        class Dev:
            capacity = n

            def renderable_extents(self):
                return ext[rank]

        class Gather:
            receives = rank == 0
            result = dict(counts=counts_tab, view_totals=counts_tab.sum(axis=0), device_ms=0.0,
                          extents=np.array([-world, -2, -3, 4, 4 + world, 6], np.float32))

            def global_list(self, v):
                g = np.concatenate([lists[r][v] + r * n for r in range(world)]).astype(np.int32)
                if corrupt and v == 1 and len(lists[0][1]) > 3:
                    g[1] += 1  # inside rank 0's own segment, which rank 0 compares entry for entry
                return torch.from_numpy(g)

        try:
            info = bench.check_exchange(torch, dist, Dev(), Gather(), "peer-root", counts_tab[rank], rank * n, nv, rank, world,
                                        device="cpu", local_list=lambda v, c: torch.from_numpy(lists[rank][v][:c]))
            q.put((rank, "ok", info["global_visible_per_view"]))
        except SystemExit as e:
            q.put((rank, "refused", str(e)))
        dist.destroy_process_group()
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, "error", traceback.format_exc()))


@pytest.mark.parametrize("corrupt", [False, True])
def test_bench_exchange_check_accepts_the_protocol_and_refuses_a_wrong_list(corrupt):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_check_worker, args=(r, 2, port, q, corrupt)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    assert all(r[1] != "error" for r in res), res
    assert all(r[1] == ("refused" if corrupt else "ok") for r in res), res
