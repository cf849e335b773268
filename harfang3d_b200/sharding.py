"""Multi-GPU sharding of batched scene instances and the visibility gather.

Scene instances are independent units (separate hg::Scene objects share nothing but read-only models), so a batch of B instances is
split into contiguous blocks, one per rank, and update + cull run with NO data-path collective. The only exchange of the path is
the gather of its result:

  1. all-gather of the per-instance visible counts (u32 per instance) — fixes every instance's offset in the global list;
  2. variable-length gather of the surviving display-entry indices to one rank (or to all ranks) with grouped point-to-point
     send/recv sized from step 1 (ncclSend/ncclRecv inside one group on NCCL; isend/irecv on gloo).

Everything here is plumbing over torch.distributed: it works on CUDA tensors with the "nccl" backend (views of the context's own
device buffers, no staging copy) and on CPU tensors with "gloo" (the world_size-2 tests). The result equals what ONE context
holding the whole batch returns (hgb_get_instance_offsets / hgb_get_visible): per-instance offsets and the concatenated visible
lists in instance order, entry indices global (local index + the rank's first entry).
"""
import numpy as np
import torch
import torch.distributed as dist


def partition_instances(n_instances, world):
    """Contiguous blocks, sizes differing by at most one: -> int64 array of world+1 instance boundaries."""
    base, rem = divmod(int(n_instances), int(world))
    sizes = np.full(world, base, np.int64)
    sizes[:rem] += 1
    return np.concatenate([[0], np.cumsum(sizes)])


def device_view(ptr, n, dtype=torch.int32, device=None):
    """Torch view (no copy) of `n` 32-bit words of a context device buffer (hgb_device_ptr)."""
    class _Arr:
        __cuda_array_interface__ = {"shape": (int(n),), "typestr": "<i4", "data": (int(ptr), False), "version": 2}
    t = torch.as_tensor(_Arr(), device=device if device is not None else "cuda")
    return t.view(dtype) if dtype != torch.int32 else t


class GatherResult:
    """Result of one gather. counts_padded: (world, pad) int32 tensor on the gather's device, row r = the visible counts of rank r's
    instances (zero padded); rank_totals: visible entries per rank (host); idx: concatenated visible entry indices in instance order
    (None on ranks that do not receive). counts / offsets (per GLOBAL instance) are derived on demand — the per-frame path never
    touches per-instance data on the host."""

    def __init__(self, counts_padded, sizes, idx, rank_totals):
        self.counts_padded, self.sizes, self.idx, self.rank_totals = counts_padded, sizes, idx, rank_totals
        self.total = int(rank_totals.sum())

    @property
    def counts(self):
        c = self.counts_padded.cpu().numpy().astype(np.int64)
        return np.concatenate([c[r, : self.sizes[r]] for r in range(len(self.sizes))]) if len(self.sizes) else np.zeros(0, np.int64)

    @property
    def offsets(self):
        return np.concatenate([[0], np.cumsum(self.counts)])

    def instance(self, i):
        off = self.offsets
        return self.idx[int(off[i]): int(off[i + 1])]


class VisibilityGather:
    """Reusable gather plan for one (instance partition, process group).

    bounds      world+1 instance boundaries (partition_instances)
    entry_base  first global display-entry index of every rank (length world): added to the local indices on the sender, so the
                gathered list indexes the whole batch's entry table
    dst         receiving rank, or None: every rank receives the full list
    """

    def __init__(self, bounds, entry_base, group=None, dst=0):
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.bounds = np.asarray(bounds, np.int64)
        assert len(self.bounds) == self.world + 1
        self.sizes = np.diff(self.bounds)
        self.pad = int(self.sizes.max()) if self.world else 0
        self.entry_base = [int(x) for x in entry_base]
        self.dst = dst
        self._cnt_pad = self._cnt_all = self._tot_all = None

    # step 1 -------------------------------------------------------------------------------------------------------------
    def gather_counts(self, local_counts):
        """local_counts: int32 tensor, one per local instance (device or CPU). -> int32 tensor (world, pad), device-side, no host sync."""
        n_loc = int(self.sizes[self.rank])
        assert local_counts.numel() == n_loc and local_counts.dtype == torch.int32
        if self._cnt_pad is None or self._cnt_pad.device != local_counts.device:
            self._cnt_pad = torch.zeros(self.pad, dtype=torch.int32, device=local_counts.device)
            self._cnt_all = torch.zeros(self.world * self.pad, dtype=torch.int32, device=local_counts.device)
        self._cnt_pad[:n_loc].copy_(local_counts)
        dist.all_gather_into_tensor(self._cnt_all, self._cnt_pad, group=self.group)
        return self._cnt_all.view(self.world, self.pad)

    # steps 1 + 2 ---------------------------------------------------------------------------------------------------------
    def gather(self, local_offsets, local_idx, out=None):
        """local_offsets: int32 tensor of n_local+1 exclusive offsets into local_idx (hgb_get_instance_offsets layout);
        local_idx: int32 tensor with the surviving local entry indices (only the first local_offsets[-1] are read).
        out: optional preallocated int32 tensor on the receiving rank(s). One host read per call: the `world` per-rank totals that size
        the variable-length transfers; the per-instance counts stay on the device."""
        counts_dev = self.gather_counts(local_offsets[1:] - local_offsets[:-1])
        if self._tot_all is None or self._tot_all.device != local_offsets.device:
            self._tot_all = torch.zeros(self.world, dtype=torch.int32, device=local_offsets.device)
        dist.all_gather_into_tensor(self._tot_all, local_offsets[-1:].contiguous(), group=self.group)
        rank_totals = self._tot_all.cpu().numpy().astype(np.int64)  # the one host read: world words
        rank_off = np.concatenate([[0], np.cumsum(rank_totals)])
        total = int(rank_off[-1])
        mine = int(rank_totals[self.rank])
        receivers = range(self.world) if self.dst is None else [self.dst]
        i_receive = self.rank in receivers

        send = None
        if mine:
            send = local_idx[:mine] + self.entry_base[self.rank] if self.entry_base[self.rank] else local_idx[:mine]
        res = None
        if i_receive:
            res = out[:total] if out is not None else torch.empty(total, dtype=torch.int32, device=local_idx.device)
            if mine:
                res[int(rank_off[self.rank]): int(rank_off[self.rank]) + mine].copy_(send)
        ops = []
        for d in receivers:
            if d != self.rank and mine:
                ops.append(dist.P2POp(dist.isend, send, self._global(d), group=self.group))
        if i_receive:
            for r in range(self.world):
                if r != self.rank and rank_totals[r]:
                    ops.append(dist.P2POp(dist.irecv, res[int(rank_off[r]): int(rank_off[r + 1])], self._global(r), group=self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return GatherResult(counts_dev, self.sizes, res, rank_totals)

    def _global(self, r):
        return dist.get_global_rank(self.group, r) if self.group is not None else r
