"""Root-subtree sharding of a scene over the GPUs of one box (new design — the
reference is one process on one thread).

One process per GPU.  Rank r owns the contiguous entity range [first_r, last_r) that starts at a
root and is closed under parent links (the reference's hierarchy utilities already treat "up to
the next self-parented entity" as one subtree, core/put/source/ecs/ecs_utilities.h:149-164), so
stages 1-2 need no communication at all.  The exchange step per frame is:

  1. all_gather of the per-view visible counts                      (world x V u32)
  2. every rank packs its V lists back to back, entity indices rebased to global numbering
     (pmb_pack_lists, one launch), and the packed runs are all_gathered (padded to the largest)
  3. segments (rank r, view v) are laid out per view in rank order  (pmb_copy_segments, one launch)
     -> view v's global list, ascending = bit-identical to the single-GPU list
  4. renderable_extents: all_reduce(min) / all_reduce(max) of 3 + 3 floats (exact, order-free)

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


class _DeviceU32:
    """a raw device pointer as something torch.as_tensor understands"""

    def __init__(self, ptr, count, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (ptr, False), "version": 2}


class ListGather:
    """The per-frame exchange for one DeviceScene (CUDA + NCCL)."""

    def __init__(self, dev, num_views, base_index, shard_entities, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.dev, self.nv, self.base, self.group = dev, int(num_views), int(base_index), group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = torch.device("cuda", dev.device)
        self.launches = 0
        self._send = None
        self._recv = None
        self._out = None
        self.counts_all = None
        self.view_offsets = None
        self.view_totals = None
        self._counts_host = torch.empty((self.world, self.nv), dtype=torch.int64).pin_memory()
        self.shard_entities = int(shard_entities)

    def _ensure(self, row_stride, total):
        torch = self.torch
        if self._send is None or self._send.numel() < row_stride:
            cap = int(row_stride * 1.25) + 1024
            self._send = torch.empty(cap, dtype=torch.int32, device=self.device)
            self._recv = torch.empty(cap * self.world, dtype=torch.int32, device=self.device)
        if self._out is None or self._out.numel() < total:
            self._out = torch.empty(int(total * 1.25) + 1024, dtype=torch.int32, device=self.device)

    def run(self):
        """Exchange the lists of the cull that just ran on dev's stream.  Returns the number of
        global list entries.  Afterwards view v's global list is self.global_list(v)."""
        torch, dist = self.torch, self.dist
        counts_dev = torch.as_tensor(_DeviceU32(self.dev.device_counts_ptr(), self.nv), device=self.device)
        counts_all = exchange_counts(counts_dev, self.group)
        self._counts_host.copy_(counts_all, non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the one host sync of the exchange: sizes the buffers
        ca = self._counts_host.numpy()
        row_totals = ca.sum(axis=1)
        row_stride = int(row_totals.max())
        row_stride = (row_stride + 3) & ~3  # keep every rank's row 16-byte aligned
        total = int(row_totals.sum())
        self.counts_all = ca.copy()
        if row_stride == 0:
            self.view_offsets = np.zeros(self.nv, np.int64)
            self.view_totals = np.zeros(self.nv, np.int64)
            return 0
        self._ensure(row_stride, total)
        self.dev.pack_lists(ca[self.rank].astype(np.uint32), self.base, self._send.data_ptr())
        send = self._send[:row_stride]
        recv = self._recv[:row_stride * self.world]
        dist.all_gather_into_tensor(recv, send, group=self.group)
        seg, self.view_offsets, self.view_totals = unpack_segments(ca, row_stride)
        self.dev.copy_segments(recv.data_ptr(), self._out.data_ptr(), seg)
        self.launches += 2
        return total

    def global_list(self, view):
        o, c = int(self.view_offsets[view]), int(self.view_totals[view])
        return self._out[o:o + c]
