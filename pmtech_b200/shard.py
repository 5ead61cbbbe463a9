"""Root-subtree sharding of a scene over the GPUs of one box (new design — the
reference is one process on one thread).

One process per GPU.  Rank r owns the contiguous entity range [first_r, last_r) that starts at a
root and is closed under parent links (the reference's hierarchy utilities already treat "up to
the next self-parented entity" as one subtree, core/put/source/ecs/ecs_utilities.h:149-164), so
stages 1-2 need no communication at all.  The exchange step per frame is:

  1. every rank packs one fixed-size row {V list lengths, the V lists back to back with entity
     indices rebased to global numbering} on the device (pmb_pack_row, one launch, lengths never
     leave the GPU)
  2. the rows are all_gathered as world x row_stride u32 (one NCCL call on a side stream, so the
     transfer overlaps the next frame's kernels; no host synchronisation anywhere)
  3. segments (rank r, view v) are laid out per view in rank order (pmb_unpack_rows, one launch,
     offsets recomputed from the gathered headers on the device)
     -> view v's global list, ascending = bit-identical to the single-GPU list
  4. renderable_extents: all_reduce(min) / all_reduce(max) of 3 + 3 floats (exact, order-free)

A row that turns out too small is reported through the gathered headers (meta[2V]); the exchange
is then repeated with larger buffers.  The functional helpers below (exchange_counts /
exchange_packed / unpack_segments, pack_row_np / unpack_rows_np) are the same protocol on host
arrays: they are what the gloo tests run, and what the GPU tests compare the kernels against.

`torch.distributed` is the plumbing (NCCL over NVLink on the GPUs; the same code runs over gloo
with CPU tensors in tests/, where the pack / unpack kernels are replaced by the test's numpy).
"""
from __future__ import annotations

import numpy as np

from . import capi
from .synth import partition_by_root, slice_scene  # noqa: F401  (re-exported: shard planning)


def unpack_segments(counts_all: np.ndarray, row_stride: int):
    """Segment table that turns the all_gathered packed rows into per-view global lists.

    counts_all[r, v] = entries rank r holds for view v; rank r's packed row starts at
    r * row_stride.  Returns (segments, view_offsets, view_totals): view v's global list is
    dst[view_offsets[v] : view_offsets[v] + view_totals[v]]."""
    counts_all = np.asarray(counts_all, np.int64)
    world, nv = counts_all.shape
    view_totals = counts_all.sum(axis=0)
    view_offsets = np.concatenate([[0], np.cumsum(view_totals)[:-1]])
    in_row = np.concatenate([np.zeros((world, 1), np.int64), np.cumsum(counts_all, axis=1)[:, :-1]], axis=1)
    before = np.concatenate([np.zeros((1, nv), np.int64), np.cumsum(counts_all, axis=0)[:-1]], axis=0)
    seg = np.zeros(world * nv, capi.SEGMENT_DTYPE)
    r_idx, v_idx = np.divmod(np.arange(world * nv), nv)
    seg["src_offset"] = r_idx * row_stride + in_row[r_idx, v_idx]
    seg["dst_offset"] = view_offsets[v_idx] + before[r_idx, v_idx]
    seg["count"] = counts_all[r_idx, v_idx]
    seg["add"] = 0
    return seg, view_offsets, view_totals


def exchange_counts(counts_local, group=None):
    """all_gather of the u32[V] visible counts -> int64 tensor [world, V] (same device)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mine = counts_local.to(torch.int32).contiguous()  # NCCL has no u32; counts < 2^31 (sb_count is an int)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return torch.stack(out).to(torch.int64)


def exchange_packed(packed_local, row_stride, group=None):
    """all_gather of the packed runs, every rank padded to row_stride entries -> [world * row_stride]."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    assert packed_local.numel() == row_stride
    out = torch.empty(world * row_stride, dtype=packed_local.dtype, device=packed_local.device)
    dist.all_gather(list(out.view(world, row_stride).unbind(0)), packed_local, group=group)
    return out


def reduce_extents(extents_local, group=None):
    """scene->renderable_extents over all shards: min of the mins, max of the maxes (exact)."""
    import torch.distributed as dist
    mn = extents_local[:3].clone()
    mx = extents_local[3:].clone()
    dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    return mn, mx


def pack_row_np(lists, add, row_stride):
    """numpy statement of pmb_pack_row: {counts[V], lists back to back (+add)}, truncated to row_stride."""
    nv = len(lists)
    row = np.zeros(row_stride, np.uint32)
    row[:nv] = [len(l) for l in lists]
    body = np.concatenate([np.asarray(l, np.uint32) + np.uint32(add) for l in lists]) if nv else np.zeros(0, np.uint32)
    m = min(body.shape[0], max(row_stride - nv, 0))
    row[nv:nv + m] = body[:m]
    return row


def unpack_rows_np(rows, nv):
    """numpy statement of pmb_unpack_rows.  rows: [world, row_stride] u32.  Returns (lists, overflow)."""
    rows = np.asarray(rows, np.uint32)
    world, row_stride = rows.shape
    counts = rows[:, :nv].astype(np.int64)
    if (counts.sum(axis=1) + nv > row_stride).any():
        return None, True
    starts = nv + np.concatenate([np.zeros((world, 1), np.int64), np.cumsum(counts, axis=1)[:, :-1]], axis=1)
    out = []
    for v in range(nv):
        out.append(np.concatenate([rows[r, starts[r, v]:starts[r, v] + counts[r, v]] for r in range(world)]))
    return out, False


class _DeviceU32:
    """a raw device pointer as something torch.as_tensor understands"""

    def __init__(self, ptr, count, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (ptr, False), "version": 2}


class ListGather:
    """The per-frame exchange for one DeviceScene (CUDA + NCCL), pipelined.

    launch() queues the pack on the scene's stream and the all-gather + unpack on a side stream and
    returns at once: the next frame's kernels overlap the transfer.  finish() waits for the last
    launch and publishes its lists (global_list).  Two sets of buffers alternate between launches."""

    def __init__(self, dev, num_views, base_index, shard_entities, group=None, scene_stream=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.dev, self.nv, self.base, self.group = dev, int(num_views), int(base_index), group
        assert self.nv <= 16, "one row carries at most PMB_MAX_VIEWS lists"
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = torch.device("cuda", dev.device)
        self.scene_stream = scene_stream if scene_stream is not None else torch.cuda.current_stream(self.device)
        self.comm = torch.cuda.Stream(self.device)
        self.launches = 0
        self.row_stride = 0
        self.shard_entities = int(shard_entities)
        self.overflows = 0
        self._k = 0
        self._send = [None, None]
        self._recv = None
        self._out = [None, None]
        self._meta = [torch.zeros(2 * 16 + 1, dtype=torch.int32, device=self.device) for _ in range(2)]
        self._meta_host = [torch.zeros(2 * 16 + 1, dtype=torch.int32).pin_memory() for _ in range(2)]
        self._packed = [torch.cuda.Event(), torch.cuda.Event()]
        self._done = [torch.cuda.Event(), torch.cuda.Event()]
        self._last = None
        self.view_offsets = None
        self.view_totals = None

    def _size(self, row_stride):
        torch = self.torch
        self.row_stride = (int(row_stride) + 3) & ~3  # keep every rank's row 16-byte aligned
        self._send = [torch.empty(self.row_stride, dtype=torch.int32, device=self.device) for _ in range(2)]
        self._recv = torch.empty(self.row_stride * self.world, dtype=torch.int32, device=self.device)
        self._out = [torch.empty(self.row_stride * self.world, dtype=torch.int32, device=self.device) for _ in range(2)]

    def _first_size(self):
        """One-time sizing from the current counts (the only host read of counts the exchange ever does)."""
        torch, dist = self.torch, self.dist
        counts_dev = torch.as_tensor(_DeviceU32(self.dev.device_counts_ptr(), self.nv), device=self.device)
        with torch.cuda.stream(self.scene_stream):
            t = counts_dev.to(torch.int64).sum().reshape(1)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        need = int(t.item())
        self._size(int(need * 1.05) + self.nv + 4096)  # lists move a little from frame to frame; an overflow is caught and repeated

    def launch(self):
        """Queue the exchange of the cull that just ran on the scene's stream."""
        torch, dist = self.torch, self.dist
        if self.row_stride == 0:
            self._first_size()
        b = self._k & 1
        self._k += 1
        with torch.cuda.stream(self.scene_stream):
            self.scene_stream.wait_event(self._done[b])  # the launch that last used this buffer set has finished
            self.dev.pack_row(self.nv, self.base, self._send[b].data_ptr(), self.row_stride)
            self._packed[b].record(self.scene_stream)
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(self._packed[b])
            dist.all_gather_into_tensor(self._recv, self._send[b], group=self.group)
            self.dev.unpack_rows(self._recv.data_ptr(), self.world, self.row_stride, self.nv, self._out[b].data_ptr(),
                                 self._out[b].numel(), self._meta[b].data_ptr(), self.comm.cuda_stream)
            self._meta_host[b].copy_(self._meta[b], non_blocking=True)
            self._done[b].record(self.comm)
        self.launches += 2
        self._last = b

    def finish(self):
        """Wait for the last launch; repeat it with larger buffers if a row overflowed.  Returns the
        number of global list entries."""
        if self._last is None:
            return 0
        for _ in range(8):
            b = self._last
            self._done[b].synchronize()
            m = self._meta_host[b].numpy()
            if not m[2 * self.nv]:
                self.view_offsets = m[:self.nv].astype(np.int64)
                self.view_totals = m[self.nv:2 * self.nv].astype(np.int64)
                return int(self.view_totals.sum())
            self.overflows += 1
            self.torch.cuda.synchronize(self.device)
            self._size(self.row_stride * 2)
            self.launch()
        raise RuntimeError("list exchange: rows still overflow after growing 8 times")

    def run(self):
        """launch() + finish(): the synchronous form."""
        self.launch()
        return self.finish()

    receives = True  # NCCL's all-gather replicates the lists

    def wait_on(self, stream):
        """make `stream` wait for everything the exchange has queued"""
        stream.wait_stream(self.comm)

    def global_list(self, view):
        o, c = int(self.view_offsets[view]), int(self.view_totals[view])
        return self._out[self._last][o:o + c]


def _result_dict(res):
    nv, world = int(res.num_views), int(res.world)
    return dict(overflow=bool(res.overflow), epoch=int(res.epoch),
                view_offsets=np.array(res.view_offset[:nv], np.int64), view_totals=np.array(res.view_total[:nv], np.int64),
                counts=np.array([list(res.counts[r][:nv]) for r in range(world)], np.int64),
                extents=np.array(res.renderable_extents[:], np.float32), device_ms=float(res.device_ms),
                lists_ptr=int(res.lists_device or 0))


class Exchange:
    """pmb_exchange for one DeviceScene: counts table + renderable extents to every rank, draw lists gathered on `root`
    (mode "root", default), replicated ("all") or left sharded ("counts"), all by NVLink peer stores from the library's
    own kernels; the cross-GPU barrier is made of device flags.  torch.distributed is used ONCE, to swap the 64-byte
    cudaIpc handles of the ranks' blocks; after that nothing of the exchange touches torch or NCCL.
    Same interface as ListGather (launch / finish / run / global_list)."""

    MODES = {"root": capi.XCHG_ROOT, "all": capi.XCHG_ALL, "counts": capi.XCHG_COUNTS}

    def __init__(self, dev, num_views, base_index, shard_entities, group=None, scene_stream=None, mode="root", root=0,
                 row_capacity=None, list_capacity=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self._C = torch, dist, C
        self.dev, self.nv, self.base, self.group = dev, int(num_views), int(base_index), group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.mode, self.root = mode, int(root)
        self.device = torch.device("cuda", dev.device)
        self.shard_entities = int(shard_entities)
        self.overflows = 0
        self.launches = 0
        self._L = capi.load()
        self._x = None
        self.row_stride = 0
        self.result = None
        self._row_capacity, self._list_capacity = row_capacity, list_capacity

    def _create(self, row_capacity, list_capacity):
        C, torch, dist = self._C, self.torch, self.dist
        self.close()
        x = C.c_void_p()
        capi._check(self._L.pmb_exchange_create(C.byref(x), self.dev._h, self.rank, self.world, self.nv, self.base, int(row_capacity),
                                                int(list_capacity), self.MODES[self.mode], self.root))
        self._x = x
        self.row_stride, self.cap = int(row_capacity), int(list_capacity)
        if self.world > 1:
            mine = np.zeros(capi.IPC_HANDLE_BYTES, np.uint8)
            capi._check(self._L.pmb_exchange_ipc_handle(x, capi._ptr(mine)))
            t = torch.from_numpy(mine).to(self.device)
            out = torch.empty(self.world * capi.IPC_HANDLE_BYTES, dtype=torch.uint8, device=self.device)
            dist.all_gather_into_tensor(out, t, group=self.group)
            handles = np.ascontiguousarray(out.cpu().numpy())
            capi._check(self._L.pmb_exchange_connect_ipc(x, capi._ptr(handles)))
            dist.barrier(group=self.group)  # every rank has opened every block before anybody stores into one

    def _first_size(self):
        """One-time sizing from the current counts: the largest row of any rank (+5 %), the global lists = world rows."""
        torch, dist = self.torch, self.dist
        if self._row_capacity is None:
            need = torch.tensor([int(self.dev.counts(self.nv).sum())], device=self.device, dtype=torch.int64)
            dist.all_reduce(need, op=dist.ReduceOp.MAX, group=self.group)
            self._row_capacity = int(int(need.item()) * 1.05) + self.nv + 4096
        if self._list_capacity is None:
            self._list_capacity = self._row_capacity * self.world
        self._create(self._row_capacity, self._list_capacity)

    def launch(self):
        if self._x is None:
            self._first_size()
        capi._check(self._L.pmb_exchange_launch(self._x))
        self.launches = int(self._L.pmb_exchange_launches(self._x))

    def finish(self):
        if self._x is None:
            return 0
        for _ in range(8):
            res = capi.ExchangeResult()
            capi._check(self._L.pmb_exchange_finish(self._x, self._C.byref(res)))
            r = _result_dict(res)
            if not r["overflow"]:
                self.result = r
                self.view_offsets, self.view_totals = r["view_offsets"], r["view_totals"]
                return int(self.view_totals.sum())
            self.overflows += 1
            self.torch.cuda.synchronize(self.device)
            self._row_capacity *= 2
            self._list_capacity = self._row_capacity * self.world
            self._create(self._row_capacity, self._list_capacity)
            self.launch()
        raise RuntimeError("list exchange: rows still overflow after growing 8 times")

    def run(self):
        self.launch()
        return self.finish()

    @property
    def receives(self):
        return self.mode == "all" or (self.mode == "root" and self.rank == self.root)

    def wait_on(self, stream):
        """enqueue what is left of the last launch, then make `stream` (a torch stream) wait for the exchange's own stream"""
        capi._check(self._L.pmb_exchange_flush(self._x))
        ext = self.torch.cuda.ExternalStream(int(self._L.pmb_exchange_stream(self._x)), device=self.device)
        stream.wait_stream(ext)

    def global_list(self, view):
        """torch int32 view of view `view`'s global list (receiving ranks only)"""
        if not self.result["lists_ptr"]:
            raise RuntimeError("this rank does not hold the global lists")
        o, c = int(self.view_offsets[view]), int(self.view_totals[view])
        return self.torch.as_tensor(_DeviceU32(self.result["lists_ptr"] + 4 * o, c), device=self.device) if c else \
            self.torch.empty(0, dtype=self.torch.int32, device=self.device)

    def close(self):
        if getattr(self, "_x", None):
            self._L.pmb_exchange_destroy(self._x)
            self._x = None

    __del__ = close


class DeviceGroup:
    """pmb_group: several GPUs (or several shards on one GPU, which is how the one-GPU test box runs the protocol) driven
    by ONE process: one pmb_scene and one exchange per rank, peer access between the devices, no torch, no NCCL."""

    def __init__(self, devices, capacities):
        import ctypes as C
        from .scene import DeviceScene
        self._C, self._L = C, capi.load()
        dv = np.ascontiguousarray(devices, np.int32)
        cp = np.ascontiguousarray(capacities, np.uint32)
        g = C.c_void_p()
        capi._check(self._L.pmb_group_create(C.byref(g), capi._ptr(dv), len(dv), capi._ptr(cp)))
        self._g = g
        self.world = len(dv)
        self.scenes = []
        class _Borrowed(DeviceScene):  # a view of one of the group's handles: the group destroys it
            def close(self):
                self._h = None

            __del__ = close

        for r in range(self.world):
            sc = _Borrowed.__new__(_Borrowed)
            sc._L, sc._h, sc.capacity, sc.device = self._L, C.c_void_p(self._L.pmb_group_scene(g, r)), int(cp[r]), int(dv[r])
            self.scenes.append(sc)

    def exchange_setup(self, num_views, base_index, row_capacity, list_capacity, mode="root", root=0):
        b = np.ascontiguousarray(base_index, np.uint32)
        capi._check(self._L.pmb_group_exchange_setup(self._g, int(num_views), capi._ptr(b), int(row_capacity), int(list_capacity),
                                                     Exchange.MODES[mode], int(root)))

    def frame(self, num_entities, frusta, mode="aabb"):
        from .scene import _MODES
        n = np.ascontiguousarray(num_entities, np.uint32)
        fr = np.ascontiguousarray(frusta, np.float32).reshape(-1, capi.FRUSTUM_FLOATS)
        capi._check(self._L.pmb_group_frame(self._g, capi._ptr(n), capi._ptr(fr), fr.shape[0], _MODES[mode]))

    def finish(self, rank=0):
        res = capi.ExchangeResult()
        capi._check(self._L.pmb_group_finish(self._g, int(rank), self._C.byref(res)))
        return _result_dict(res)

    def close(self):
        if getattr(self, "_g", None):
            self._L.pmb_group_destroy(self._g)
            self._g = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def make_gather(dev, num_views, base_index, shard_entities, group=None, scene_stream=None, prefer="peer", mode="root"):
    """The library's own peer-memory exchange (Exchange) when the GPUs of the box can map each other's memory, else the
    NCCL ListGather (which replicates the lists).  Returns (gather, kind)."""
    import torch
    import torch.distributed as dist
    if prefer == "peer":
        ok = 1
        g = None
        try:
            g = Exchange(dev, num_views, base_index, shard_entities, group, scene_stream, mode=mode)
            g._first_size()
        except Exception as e:  # no peer mapping on this box / driver
            import sys
            print(f"[pmtech_b200.shard] peer-memory exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
            ok = 0
        t = torch.tensor([ok], device=torch.device("cuda", dev.device))
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
        if int(t.item()) == 1:
            return g, "peer-" + mode
        if g is not None:
            g.close()
    return ListGather(dev, num_views, base_index, shard_entities, group, scene_stream), "nccl"
