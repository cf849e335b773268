"""Multi-rank visibility gather on CPU: world_size 2 and 3 over gloo, 127.0.0.1.

Every rank owns a contiguous block of the instance batch (sharding.partition_instances), computes ITS instances' visible lists — here
with the oracle, because there is no GPU in this container; on the GPU box tests/test_gpu_parity.py feeds the same gather from the
CUDA path — and the gather must reproduce exactly what one scene holding the whole batch gives: per-instance counts, offsets and the
concatenated lists with global entry indices. Edge cases: ranks with zero visible entries, more ranks than instances, dst=None
(every rank receives)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from harfang3d_b200 import hostmirror, scenegen, sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _instance_lists(fs):
    """oracle: per-instance visible entry indices (into fs's kind-0 entry table) and the entry table size"""
    from tests import oracle_lib
    from tests.port_scene import PortScene
    t = hostmirror.build_tables(fs)
    orc = PortScene(oracle_lib.load_port(), fs, t)
    orc.update(0)
    frusta = hostmirror.scene_frusta(fs)
    flags = t["ent0_flags"]
    enabled = np.nonzero(flags == 0)[0]
    ent_node = t["ent0_mtx_slot"]
    lists = []
    for i in range(len(frusta)):
        pos_ref, _ = orc.cull(0, frusta[i])
        idx = enabled[pos_ref]
        lo, hi = fs.instance_start[i], fs.instance_start[i + 1]
        lists.append(idx[(ent_node[idx] >= lo) & (ent_node[idx] < hi)].astype(np.int32))
    return lists, len(ent_node)


def _worker(rank, world, port, n_inst, dst, blank_rank, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fs = scenegen.config_c5(seed=11, n_instances=n_inst, nodes_per_instance=192, depth=4)
        if blank_rank is not None:  # one rank sees nothing: all of its nodes disabled
            b = sharding.partition_instances(n_inst, world)
            fs.disabled[fs.instance_start[b[blank_rank]]:fs.instance_start[b[blank_rank + 1]]] = 1
        bounds = sharding.partition_instances(n_inst, world)
        i0, i1 = int(bounds[rank]), int(bounds[rank + 1])
        # entry_base of every rank = number of kind-0 entries owned by the ranks before it (every node has one entry here)
        whole_t = hostmirror.build_tables(fs)
        ent_node = whole_t["ent0_mtx_slot"]
        entry_base = [int(np.searchsorted(ent_node, fs.instance_start[bounds[r]])) for r in range(world)]
        if i1 > i0:
            sub = scenegen.slice_instances(fs, i0, i1)
            lists, _ = _instance_lists(sub)
        else:
            lists = []
        off = np.concatenate([[0], np.cumsum([len(x) for x in lists])]).astype(np.int32)
        idx = np.concatenate(lists + [np.zeros(0, np.int32)]).astype(np.int32)
        g = sharding.VisibilityGather(bounds, entry_base, dst=dst)
        # the local list buffer is larger than the survivors (capacity = entry table), like the context's device buffer
        res = g.gather(torch.from_numpy(off), torch.from_numpy(np.concatenate([idx, np.full(7, -1, np.int32)])))
        receives = dst is None or rank == dst
        assert (res.idx is not None) == receives
        if receives:
            q.put((rank, res.counts.tolist(), res.offsets.tolist(), res.idx.numpy().tolist()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_inst,dst,blank", [(2, 7, 0, None), (2, 4, None, 1), (3, 2, 0, None), (3, 8, 2, 0)])
def test_gather_equals_whole_batch(world, n_inst, dst, blank):
    fs = scenegen.config_c5(seed=11, n_instances=n_inst, nodes_per_instance=192, depth=4)
    if blank is not None:
        b = sharding.partition_instances(n_inst, world)
        fs.disabled[fs.instance_start[b[blank]]:fs.instance_start[b[blank + 1]]] = 1
    lists, _ = _instance_lists(fs)
    want_counts = [len(x) for x in lists]
    want_idx = np.concatenate(lists).tolist()
    assert sum(want_counts) > 0

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_inst, dst, blank, q)) for r in range(world)]
    [p.start() for p in procs]
    got = [q.get(timeout=180) for _ in range(world if dst is None else 1)]
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, counts, offsets, idx in got:
        assert counts == want_counts, "rank %d" % rank
        assert offsets == np.concatenate([[0], np.cumsum(want_counts)]).tolist()
        assert idx == want_idx, "rank %d" % rank


def test_partition_instances():
    assert sharding.partition_instances(8192, 8).tolist() == [1024 * i for i in range(9)]
    assert sharding.partition_instances(7, 2).tolist() == [0, 4, 7]
    assert sharding.partition_instances(2, 3).tolist() == [0, 1, 2, 2]
    b = sharding.partition_instances(1000, 7)
    assert b[0] == 0 and b[-1] == 1000 and np.diff(b).max() - np.diff(b).min() <= 1


# ---- peer-gather setup: the host-side handshake (capi.Context.peer_setup) must leave every rank the same way ----
class _FakePeerContext:
    """the device calls of the handshake stubbed out (no GPU here); peer_setup itself is the product's"""

    def __init__(self, rank, fail_init_on, fail_connect_on):
        self.rank, self.fail_init_on, self.fail_connect_on = rank, fail_init_on, fail_connect_on
        self.connected = self.closed = False

    def peer_init(self, rank, world, dst, entries_per_rank, aux_words_per_rank=None):
        from harfang3d_b200 import capi
        if rank == self.fail_init_on:
            raise capi.HgbError("no IPC here")
        return bytes([rank]) * 64, bytes([100 + rank]) * 64

    def peer_connect(self, inbox_of_dst, mailboxes):
        from harfang3d_b200 import capi
        assert inbox_of_dst == bytes([0]) * 64 and [m[0] for m in mailboxes] == [100 + r for r in range(len(mailboxes))]
        if self.rank == self.fail_connect_on:
            raise capi.HgbError("no peer access")
        self.connected = True

    def peer_close(self):
        self.closed, self.connected = True, False


def _peer_worker(rank, world, port, fail_init_on, fail_connect_on, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from harfang3d_b200 import capi
        c = _FakePeerContext(rank, fail_init_on, fail_connect_on)
        try:
            capi.Context.peer_setup(c, dist, np.full(world, 10, np.uint32), None, dst=0)
            q.put((rank, "ok", c.connected, c.closed))
        except capi.HgbError as e:
            q.put((rank, str(e), c.connected, c.closed))
        dist.barrier()  # every rank is still in step with the others
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,fail_init_on,fail_connect_on", [(2, None, None), (2, 1, None), (3, None, 2), (3, 0, None)])
def test_peer_setup_fails_on_every_rank_or_none(world, fail_init_on, fail_connect_on):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_worker, args=(r, world, port, fail_init_on, fail_connect_on, q)) for r in range(world)]
    [p.start() for p in procs]
    got = sorted(q.get(timeout=120) for _ in range(world))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    if fail_init_on is None and fail_connect_on is None:
        assert all(g[1] == "ok" and g[2] and not g[3] for g in got)
    else:
        what = "no IPC here" if fail_init_on is not None else "no peer access"
        assert all(what in g[1] and "peer gather unavailable" in g[1] and not g[2] and g[3] for g in got), got
