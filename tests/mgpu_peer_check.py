"""Run under torchrun (one process per GPU): the peer-memory visibility gather (hgb_peer_*) against ONE context that holds the whole
instance batch. Every rank builds the same seeded batch, keeps its contiguous block of instances (sharding.partition_instances),
and runs fused frames; the destination rank's concatenated list, rank offsets and global per-instance offsets must equal
hgb_get_visible / hgb_get_instance_offsets of the single context, bit for bit. Also exercises the flow control (frames enqueued
back to back without host synchronisation) and a shared-camera batch (no instance mode, no aux words).
Started by tests/test_gpu_multi.py when the box has at least 2 GPUs, or by scripts/gpu_multi.sh."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import harfang3d_b200  # noqa: E402
from harfang3d_b200 import hostmirror, scenegen, sharding  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    # ---- instance batch, one camera per instance ----
    n_inst = 61  # not a multiple of the world size: ragged blocks
    fs = scenegen.config_c5(seed=77, n_instances=n_inst, nodes_per_instance=700, depth=5)
    bounds = sharding.partition_instances(n_inst, world)
    mine = scenegen.slice_instances(fs, int(bounds[rank]), int(bounds[rank + 1]))
    ctx = harfang3d_b200.Context(local)
    t = hostmirror.upload(ctx, mine)
    ctx.update(None, fused_cull=True)  # builds the device tables
    ctx.sync()
    ent = torch.zeros(world, dtype=torch.int64, device="cuda")
    ent[rank] = len(t["ent0_mtx_slot"])
    dist.all_reduce(ent)
    ent = ent.cpu().numpy()
    ctx.peer_setup(dist, ent, np.diff(bounds), dst=0)
    frames = 25
    for f in range(frames):  # back to back: the ack words are the only thing that keeps a fast rank from overwriting an unread half
        if f % 5 == 0:       # move something so that the lists differ from frame to frame
            trs = mine.trs.copy()
            trs[:, 0] += 0.37 * (f + 1)
            ctx.upload_trs(np.ascontiguousarray(trs, np.float32))
        ctx.update(None, fused_cull=True)
    ctx.sync()
    step, err = ctx.peer_status()
    assert err == 0, "rank %d: peer gather error word %d" % (rank, err)
    assert step == frames, (step, frames)
    dist.barrier()
    ok = 1
    if rank == 0:
        off, idx, aux = ctx.peer_result()
        whole = harfang3d_b200.Context(local)
        full = scenegen.config_c5(seed=77, n_instances=n_inst, nodes_per_instance=700, depth=5)
        full.trs[:, 0] += np.float32(0.37 * 21)  # the state after the last upload (f = 20), applied to every rank's block alike
        hostmirror.upload(whole, full)
        whole.update(None, fused_cull=True)
        idx_ref, off_ref = whole.get_visible(0, 0), whole.get_instance_offsets()
        assert len(idx) == len(idx_ref) and (idx == idx_ref).all(), "gathered list differs from the single-context list"
        assert (aux == off_ref).all(), "global per-instance offsets differ"
        assert off[0] == 0 and off[-1] == len(idx_ref) and (np.diff(off.astype(np.int64)) >= 0).all()
        for r in range(world):
            assert off[r] == off_ref[int(bounds[r])]
        whole.close()
        print("peer gather (instances): %d ranks, %d instances, %d visible entries, %d frames: identical to one context" % (world, n_inst, len(idx), frames))
    ctx.close()

    # ---- shared camera (c4 style): every rank its own scenes, no aux words ----
    fs2 = scenegen.static_scene(30_000, 6, 6, n_models=3, seed=500 + rank)
    ctx = harfang3d_b200.Context(local)
    t = hostmirror.upload(ctx, fs2)
    ctx.update(None, fused_cull=True)
    ctx.sync()
    ent = torch.zeros(world, dtype=torch.int64, device="cuda")
    ent[rank] = len(t["ent0_mtx_slot"])
    dist.all_reduce(ent)
    ent = ent.cpu().numpy()
    ctx.peer_setup(dist, ent, None, dst=0)
    for _ in range(7):
        ctx.update(None, fused_cull=True)
    ctx.sync()
    assert ctx.peer_status()[1] == 0
    local_idx = torch.from_numpy(ctx.get_visible(0, 0).astype(np.int64)).cuda()
    n_loc = torch.tensor([local_idx.numel()], device="cuda")
    counts = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
    dist.all_gather(counts, n_loc)
    counts = [int(c.item()) for c in counts]
    pad = torch.zeros(max(counts), dtype=torch.int64, device="cuda")
    pad[: local_idx.numel()] = local_idx
    allp = [torch.zeros_like(pad) for _ in range(world)]
    dist.all_gather(allp, pad)
    if rank == 0:
        off, idx, _ = ctx.peer_result()
        base = np.concatenate([[0], np.cumsum(ent)[:-1]])
        want = np.concatenate([allp[r][: counts[r]].cpu().numpy() + base[r] for r in range(world)]).astype(np.uint32)
        assert (off == np.concatenate([[0], np.cumsum(counts)])).all()
        assert (idx == want).all()
        print("peer gather (shared camera): %d ranks, %d visible entries: identical to the per-rank lists" % (world, len(idx)))

    # ---- asynchronous tail + skewed host submission + a rank WITHOUT display entries ----
    # (the gather's step counter lives in the tail's own stream-ordered sequence: ranks that run ahead or behind on the host, and a rank
    # that publishes an empty list, must all land in the same step)
    import time
    dist.barrier()
    ctx.peer_close()
    ctx.close()
    fs3 = scenegen.static_scene(40_000, 6, 6, n_models=3, seed=900 + rank)
    if rank == world - 1:
        fs3.model[:] = -1  # no object anywhere: zero display entries on this rank
    ctx = harfang3d_b200.Context(local, async_post=True)
    t = hostmirror.upload(ctx, fs3)
    ctx.update(None, fused_cull=True)
    ctx.sync()
    ent = torch.zeros(world, dtype=torch.int64, device="cuda")
    ent[rank] = len(t["ent0_mtx_slot"])
    dist.all_reduce(ent)
    ent = ent.cpu().numpy()
    assert ent[world - 1] == 0
    ctx.peer_setup(dist, ent, None, dst=0)
    for f in range(16):
        if (f + rank) % 3 == 0:
            time.sleep(0.003 * (1 + rank % 3))  # host-side skew between the ranks
        if f % 4 == 1:
            trs = fs3.trs.copy()
            trs[:, 2] += 0.21 * (f + 1)
            ctx.upload_trs(np.ascontiguousarray(trs, np.float32))
        ctx.update(None, fused_cull=True)
    ctx.sync()
    st, err = ctx.peer_status()
    assert err == 0 and st == 16, (st, err)
    local_idx = torch.from_numpy(ctx.get_visible(0, 0).astype(np.int64)).cuda()
    n_loc = torch.tensor([local_idx.numel()], device="cuda")
    counts = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
    dist.all_gather(counts, n_loc)
    counts = [int(c.item()) for c in counts]
    pad = torch.zeros(max(max(counts), 1), dtype=torch.int64, device="cuda")
    pad[: local_idx.numel()] = local_idx
    allp = [torch.zeros_like(pad) for _ in range(world)]
    dist.all_gather(allp, pad)
    if rank == 0:
        off, idx, _ = ctx.peer_result()
        base = np.concatenate([[0], np.cumsum(ent)[:-1]])
        want = np.concatenate([allp[r][: counts[r]].cpu().numpy() + base[r] for r in range(world)]).astype(np.uint32)
        assert counts[world - 1] == 0
        assert (off == np.concatenate([[0], np.cumsum(counts)])).all(), (off, counts)
        assert len(idx) == len(want) and (idx == want).all()
        print("peer gather (asynchronous tail, skewed hosts, one empty rank): %d ranks, %d visible entries, 16 frames: identical to the per-rank lists" % (world, len(idx)))
    ent_keep = ent

    # ---- close, then a setup that fails on ONE rank: every rank must leave peer_setup with the same error and keep running frames ----
    before = ctx.get_visible(0, 0).copy()
    dist.barrier()
    ctx.peer_close()
    ctx.update(None, fused_cull=True)
    assert (ctx.get_visible(0, 0) == before).all()
    if rank == world - 1:
        def broken(*a, **k):
            raise harfang3d_b200.capi.HgbError("injected: no IPC on this rank")
        ctx.peer_init = broken
    try:
        ctx.peer_setup(dist, ent, None, dst=0)
        raise AssertionError("peer_setup succeeded with a broken rank")
    except harfang3d_b200.capi.HgbError as e:
        assert "injected" in str(e), e
    ctx.update(None, fused_cull=True)
    assert (ctx.get_visible(0, 0) == before).all()
    if rank == 0:
        print("peer close + one-rank setup failure: every rank fell back together, frames unchanged")
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    print("rank %d ok" % rank)


if __name__ == "__main__":
    main()
