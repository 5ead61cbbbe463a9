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

    def global_list(self, view):
        o, c = int(self.view_offsets[view]), int(self.view_totals[view])
        return self._out[self._last][o:o + c]


class PeerGather(ListGather):
    """The same exchange as ONE push over peer memory instead of all-gather + unpack: every rank's
    final per-view lists live in a symmetric buffer (torch.distributed._symmetric_memory, NVLink P2P
    mappings of the same allocation on every GPU of the box); after the list lengths have been swapped
    by P2P stores (pmb_publish_counts + barrier), pmb_push_row writes each local segment straight to
    its final place in every peer's lists (+ barrier).  Nothing is received into an intermediate
    buffer and nothing is unpacked.  Same interface as ListGather; raises at construction when the
    symmetric-memory rendezvous is not available (use make_gather for the NCCL fallback)."""

    TAB = 16 * 16  # u32[PMB_MAX_PEERS][PMB_MAX_VIEWS]

    def __init__(self, dev, num_views, base_index, shard_entities, group=None, scene_stream=None):
        super().__init__(dev, num_views, base_index, shard_entities, group, scene_stream)
        import torch.distributed._symmetric_memory as symm
        self._symm = symm
        self._hdl = None
        self._buf = None

    def _size(self, row_stride):
        torch, dist = self.torch, self.dist
        self.row_stride = (int(row_stride) + 3) & ~3
        self.cap = self.row_stride * self.world  # entries of one set of global lists
        self._send = [torch.empty(self.row_stride, dtype=torch.int32, device=self.device) for _ in range(2)]
        words = 2 * self.TAB + 2 * self.cap
        torch.cuda.synchronize(self.device)
        self._buf = self._symm.empty(words, dtype=torch.int32, device=self.device)
        self._buf.zero_()
        grp = self.group if self.group is not None else dist.group.WORLD
        self._hdl = self._symm.rendezvous(self._buf, grp)
        ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        assert len(ptrs) == self.world
        self._tab_ptrs = [np.array([p + 4 * b * self.TAB for p in ptrs], np.uint64) for b in range(2)]
        self._out_ptrs = [np.array([p + 4 * (2 * self.TAB + b * self.cap) for p in ptrs], np.uint64) for b in range(2)]
        self._tab_local = [self._buf[b * self.TAB:(b + 1) * self.TAB] for b in range(2)]
        self._out = [self._buf[2 * self.TAB + b * self.cap:2 * self.TAB + (b + 1) * self.cap] for b in range(2)]
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def launch(self):
        torch = self.torch
        if self.row_stride == 0:
            self._first_size()
        b = self._k & 1
        self._k += 1
        with torch.cuda.stream(self.scene_stream):
            self.scene_stream.wait_event(self._done[b])
            self.dev.pack_row(self.nv, self.base, self._send[b].data_ptr(), self.row_stride)
            self._packed[b].record(self.scene_stream)
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(self._packed[b])
            cs = self.comm.cuda_stream
            self.dev.publish_counts(self._send[b].data_ptr(), self.nv, self.rank, self.world, self._tab_ptrs[b], cs)
            self._hdl.barrier(channel=0)  # every rank's table holds all list lengths
            self.dev.push_row(self._send[b].data_ptr(), self.row_stride, self.nv, self.rank, self.world,
                              self._tab_local[b].data_ptr(), self._out_ptrs[b], self.cap, self._meta[b].data_ptr(), cs)
            self._hdl.barrier(channel=1)  # every rank's lists are complete
            self._meta_host[b].copy_(self._meta[b], non_blocking=True)
            self._done[b].record(self.comm)
        self.launches += 3
        self._last = b


def make_gather(dev, num_views, base_index, shard_entities, group=None, scene_stream=None, prefer="peer"):
    """PeerGather when the GPUs of the box can map each other's memory, else the NCCL ListGather.
    Returns (gather, kind)."""
    import torch.distributed as dist
    if prefer == "peer":
        ok = 1
        try:
            g = PeerGather(dev, num_views, base_index, shard_entities, group, scene_stream)
            g.launch()  # sizes the buffers and runs the rendezvous
            g.finish()
        except Exception as e:  # rendezvous unsupported on this box / driver
            import sys
            print(f"[pmtech_b200.shard] peer-memory exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
            ok = 0
        import torch
        t = torch.tensor([ok], device=torch.device("cuda", dev.device))
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
        if int(t.item()) == 1:
            return g, "peer"
    return ListGather(dev, num_views, base_index, shard_entities, group, scene_stream), "nccl"
