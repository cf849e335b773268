#!/usr/bin/env python
"""bench.py — scene nodes updated+culled per second on B200 (BASELINE.json metric), one JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c5] [--impl reference]

A "step" is one frame of the hot path over one batch of synthetic scenes: Scene::Update (SceneAnim key sampling + world-matrix
propagation) fused with GetModelDisplayLists/CullModelDisplayLists (visibility + order-preserving compaction).
Workload at every N (weak scaling, per-GPU work fixed): BASELINE.json configs[3] — the 16M-node mixed static/animated scene batch
(16 scenes x 1M nodes, depth 8, 25 % of the nodes driven by a Hermite Position track) — the configuration the metric's roofline
target is quoted on; every rank owns its own batch of independent scene instances (different seed), and at N>1 the per-rank
visible counts are all-gathered over NCCL inside the step (the path has no other exchange).

  value          nodes/s with the scene resident in HBM, CUDA events on the launching stream, max over ranks
  e2e            the same metric through the C ABI with HOST buffers inside the timed region: every step uploads the TRS of EVERY
                 node from pinned host memory (hgb_upload_trs: all transforms host-edited, the physics/script-driven worst case),
                 the play clocks and frusta, and reads the visible list back to pinned host memory (hgb_get_visible)
  e2e_resident   the Scene::Update(dt) drop-in case: scene state resident (as it is inside hg::Scene on the CPU), per step only the
                 play clocks go up and the visible list comes down
  roofline       k_update_level (fused update+cull), algorithmic bytes per node / CUDA-event time per launch
  cpu_baseline   the UNMODIFIED reference (oracle/_ref) on this box's host cores, bounded sample of the same workload

--impl reference times the reference's own CPU implementation (oracle/_ref when built, else the C port) with all host threads
(one hg::Scene per thread: the reference path itself is single-threaded, scene instances are independent).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from harfang3d_b200 import hostmirror, scenegen  # noqa: E402

METRIC = "scene nodes updated+culled/sec"
UNIT = "Mnodes/s"
FRAME_DT = scenegen.FRAME_DT
# algorithmic bytes (SURVEY.md §8d / DESIGN.md §5): fused update+cull reads TRS 36 + parent 4 + entry 8 + AABB 24, writes world 48,
# plus 4 B per surviving entry
FUSED_BYTES_PER_NODE = 120


def workload(name, rank, scale=1.0):
    """-> (FlatScene, description). scale < 1 shrinks the node count (smoke runs / CPU samples)."""
    if name == "c4":
        n_scenes = max(1, int(round(16 * scale)))
        fs = scenegen.config_c4(seed=4 + 1000 * rank, n_scenes=n_scenes, nodes_per_scene=1_000_000, animated_frac=0.25, n_keys=16)
        return fs, "c4: %d scenes x 1M nodes, depth 8, 25%% animated (Position track, 16 keys), fused update+cull, 1 camera" % n_scenes
    if name == "c2":
        fs = scenegen.config_c2(seed=2 + 1000 * rank, n_nodes=max(1000, int(1_000_000 * scale)))
        return fs, "c2: %d-node static scene, depth 8, fused update+cull" % fs.n
    if name == "c5":
        # per-rank share of BASELINE configs[4] at 2 GPUs (8192 instances x 2048 nodes = 16.8M slots exceed the reference's 2^24-1
        # vector_list capacity for ONE scene/context, foundation/vector_list.h:40)
        fs = scenegen.config_c5(seed=5 + 1000 * rank, n_instances=max(1, int(4096 * scale)))
        return fs, "c5: %d instances x 2048 nodes per GPU, depth 6, one camera per instance" % (fs.n // 2048)
    raise SystemExit("unknown workload %s" % name)


# ---- clocks ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); smax.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v == "Active":
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---- CPU legs (the only place outside tests/ that may touch oracle/) -------------------------------------------------------
def cpu_reference_rate(threads, n_scenes, nodes_per_scene, frames, seed=4, warm=1):
    """nodes/s of the reference's own Scene::Update + GetModelDisplayLists + CullModelDisplayLists on `threads` host threads.
    The sample is `n_scenes` scenes of the c4 batch (depth 8, 25 % animated, one hg::Scene each — the reference path itself is
    single-threaded, scene instances are independent), dealt round-robin to the threads; a frame = every scene stepped once.
    Returns (rate, kind, description)."""
    from tests import oracle_lib
    from tests.oracle_lib import P
    ref = oracle_lib.load_ref()
    kind = "reference" if ref is not None else "port"
    threads = max(1, min(threads, n_scenes))
    scenes = [None] * n_scenes

    def build(th):
        for i in range(th, n_scenes, threads):
            fs = scenegen.config_c4(seed=seed + i, n_scenes=1, nodes_per_scene=nodes_per_scene, animated_frac=0.25, n_keys=16)
            fr = np.ascontiguousarray(hostmirror.scene_frusta(fs)[0])
            if ref is not None:
                from tests.scene_io import RefScene
                scenes[i] = (RefScene(ref, fs), fr)
            else:
                from tests.port_scene import PortScene
                scenes[i] = (PortScene(oracle_lib.load_port(), fs), fr)

    def run(th, n_frames):
        for _ in range(n_frames if ref is None else 1):
            for i in range(th, n_scenes, threads):
                sc, fr = scenes[i]
                if ref is not None:  # scenes are independent: n_frames of one scene, then the next, is the same work as frame-major order
                    out = np.zeros(4)
                    ref.hgref_time_frames(sc.h, n_frames, FRAME_DT, P(fr), 0, P(out))
                else:
                    sc.update(FRAME_DT)
                    sc.cull(0, fr)

    def par(fn, *a):
        th = [threading.Thread(target=fn, args=(i,) + a) for i in range(threads)]
        [t.start() for t in th]
        [t.join() for t in th]

    par(build)
    par(run, max(warm, 1))  # warm-up frames (the first Update sizes the vectors)
    t0 = time.perf_counter()
    par(run, frames)
    wall = time.perf_counter() - t0
    for sc, _ in scenes:
        sc.close()
    rate = n_scenes * nodes_per_scene * frames / wall
    return rate, kind, "%d hg::Scene(s) of %d nodes (scenes of the c4 batch) on %d thread(s) x %d frames, wall %.2f s" % (
        n_scenes, nodes_per_scene, threads, frames, wall)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference_arm(args, rank):
    """The reference's own CPU implementation on all host threads, on the c4 batch itself: 16 hg::Scenes of 1M nodes (scaled by --scale),
    one step = one frame of every scene."""
    if rank != 0:
        return
    threads = host_threads()
    n_scenes = max(1, int(round(16 * args.scale)))
    nodes = 1_000_000
    rate, kind, sample = cpu_reference_rate(threads, n_scenes, nodes, args.steps, warm=args.warmup)
    v = rate / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": n_scenes * nodes / rate * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "c4: %d scenes x 1M nodes, depth 8, 25%% animated (Position track, 16 keys), Scene::Update + GetModelDisplayLists + "
                                   "CullModelDisplayLists, 1 camera; CPU: " % n_scenes + sample, "scale": args.scale},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": min(threads, n_scenes), "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---- own arm ------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default="c4", choices=["c4", "c2", "c5"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; the reported config says so)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lookahead", type=int, default=0, help="A/B switch: L2 prefetch distance in positions (0 = default, -1 = off)")
    ap.add_argument("--tma-level", action="store_true", help="A/B switch: persistent TMA-staged level kernel instead of one thread per node (LDG rows)")
    ap.add_argument("--top-walk", type=int, default=0, help="A/B switch: budget (node evaluations) of the one-launch top of the hierarchy (0 = default, -1 = one launch per level)")
    ap.add_argument("--slice-nodes", type=int, default=0, help="A/B switch: slice size of big levels under the overlapped sampler (0 = no slicing)")
    ap.add_argument("--overlap", action="store_true", help="A/B switch: sample tracks by target level on a second stream, overlapped with the level kernels")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"], help="N>1 result gather: peer = survivors stored into rank 0's memory over NVLink "
                    "by the compaction kernel itself, flags + acks on the device, no host sync (hgb_peer_*); nccl = counts all-gathered, host-sized send/recv (A/B)")
    ap.add_argument("--async-post", type=int, default=None, choices=[0, 1], help="A/B switch: compaction + gather of a frame on a second stream, overlapping "
                    "the next frame's track sampling (hgb_options.async_post)")
    ap.add_argument("--row-store", action="store_true", help="A/B switch: per-thread world row stores instead of the transposed warp store")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    import harfang3d_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this path has no CPU fallback (use --impl reference for the CPU reference)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    fs, desc = workload(args.workload, rank, args.scale)
    n_nodes = fs.n
    ctx = harfang3d_b200.Context(local_rank, row_store=args.row_store, lookahead=args.lookahead, tma_level=args.tma_level, top_walk=args.top_walk, overlap_sampler=args.overlap, slice_nodes=args.slice_nodes,
                                  async_post=(False if (world > 1 and args.exchange == "nccl") else None) if args.async_post is None else bool(args.async_post))  # the NCCL A/B reads the lists on the caller's stream: no asynchronous tail there
    t = hostmirror.upload(ctx, fs)
    n_entries = len(t["ent0_mtx_slot"]) + len(t["ent1_mtx_slot"])
    clocks = hostmirror.PlayClocks(fs)
    K, W = args.steps, max(args.warmup, 3)

    # N>1: the path's only exchange — visible counts all-gathered, visible lists gathered to rank 0 with grouped
    # send/recv sized from the counts. The tensors are views of the context's own device buffers (no staging copy).
    gather = None
    gathered = [0]
    peer = world > 1 and args.exchange == "peer"
    if world > 1:
        from harfang3d_b200 import sharding
        if not peer:
            ctx.set_stream(torch.cuda.current_stream().cuda_stream)  # one stream for our kernels and the NCCL gather
        ctx.update(clocks.step(0), fused_cull=True)              # first frame: builds the device tables the views below point into
        ctx.sync()
        n_ent0 = len(t["ent0_mtx_slot"])
        ent_all = torch.zeros(world, dtype=torch.int64, device="cuda")
        ent_all[rank] = n_ent0
        dist.all_reduce(ent_all)
        ent_all = ent_all.cpu().numpy()
        entry_base = np.concatenate([[0], np.cumsum(ent_all)[:-1]])
    if peer:
        # the path's only exchange rides on the frame itself: k_compact stores the survivors into rank 0's inbox over
        # NVLink, flags and acknowledgements are device words, rank 0 concatenates — nothing of it touches the host (hgb_kernels_peer.cuh)
        aux = np.full(world, len(fs.instance_cameras), np.uint32) if fs.instance_cameras is not None else None
        try:
            ctx.peer_setup(dist, ent_all, aux, dst=0)
        except harfang3d_b200.capi.HgbError as e:  # raised on every rank or on none: the other device-side exchange takes over, and the line says so
            peer = False
            args.exchange = "nccl"
            if rank == 0:
                print("bench: %s -> --exchange nccl" % e, file=sys.stderr)
            ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    if world > 1 and not peer:
        # A/B: visible counts all-gathered, visible lists gathered to rank 0 with grouped send/recv sized from the counts (one host read
        # per step). The tensors are views of the context's own device buffers (no staging copy).
        if fs.instance_cameras is not None:  # c5: per-instance offsets of the compacted list
            n_inst = len(fs.instance_cameras)
            bounds = np.arange(world + 1, dtype=np.int64) * n_inst
            rank_view = sharding.device_view(ctx.device_ptr(5), n_inst)  # exclusive ranks of the instances' first entries
            off_view = torch.zeros(n_inst + 1, dtype=torch.int32, device="cuda")
            cnt_view = sharding.device_view(ctx.device_ptr(3), 1)
        else:                                # c4: one camera per rank batch = one "instance" per rank
            bounds = np.arange(world + 1, dtype=np.int64)
            off_view = torch.zeros(2, dtype=torch.int32, device="cuda")
            cnt_view = sharding.device_view(ctx.device_ptr(3), 1)
        idx_view = sharding.device_view(ctx.device_ptr(2), max(n_ent0, 1))
        gather_out = torch.empty(int(ent_all.sum()), dtype=torch.int32, device="cuda") if rank == 0 else None
        gather = sharding.VisibilityGather(bounds, entry_base, dst=0)

    def barrier():
        ctx.sync()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        ctx.update(clocks.step(FRAME_DT), fused_cull=True)
        if gather is not None:
            if fs.instance_cameras is not None:
                off_view[:-1].copy_(rank_view)
            off_view[-1:].copy_(cnt_view)
            res = gather.gather(off_view, idx_view, out=gather_out)
            gathered[0] = res.total

    # ---- warm-up, then the device-resident timed region -------------------------------------------------------------------
    for _ in range(W):
        step_resident()
    barrier()
    sampler = ClockSampler(local_rank)
    launches0 = ctx.launch_count
    ctx.time_begin()
    for _ in range(K):
        step_resident()
    ms = ctx.time_end()
    launches = ctx.launch_count - launches0
    barrier()
    n_visible = ctx.get_visible_count(0, 0)

    # ---- per-kernel CUDA-event times over K more steps (plain launches, event pair around every kernel) -------------------
    ctx.set_profiling(True)
    ctx.get_profile(reset=True)
    for _ in range(K):
        step_resident()
    prof = ctx.get_profile(reset=True)
    ctx.set_profiling(False)
    lvl_ms, lvl_n = prof["k_update_level"]
    per_step_ms = {k: v[0] / K for k, v in prof.items() if v[1]}

    # ---- end to end through the C ABI with host buffers ---------------------------------------------------------------------
    pin_trs = torch.empty((n_nodes, 9), dtype=torch.float32, pin_memory=True).numpy()
    pin_trs[:] = fs.trs
    pin_out = torch.empty(max(len(t["ent0_mtx_slot"]), 1), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)

    def step_e2e(full_upload):
        if full_upload:
            ctx.upload_trs(pin_trs)
        step_resident()
        return ctx.get_visible(0, 0, out=pin_out)

    e2e = {}
    for name, full in (("e2e", True), ("e2e_resident", False)):
        for _ in range(3):
            vis = step_e2e(full)
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            vis = step_e2e(full)
        ctx.sync()
        dt_s = time.perf_counter() - t0
        barrier()
        h2d = (n_nodes * 36 if full else 0) + 8 * len(fs.plays)
        e2e[name] = [dt_s, h2d, 4 + 4 * len(vis)]

    if peer:
        step_no, err = ctx.peer_status()
        if err:
            raise SystemExit("bench.py: peer gather error word %d on rank %d after %d frames" % (err, rank, step_no))
        if rank == 0:
            gathered[0] = int(ctx.peer_result(with_idx=False)[0][-1])
    clocks_rec = sampler.stop()  # sampled from the start of the timed region to the end of the e2e legs (all under load)

    # ---- reduce over ranks --------------------------------------------------------------------------------------------------
    def rmax(x):
        if world == 1:
            return x
        tt = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def rsum(x):
        if world == 1:
            return x
        tt = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        return float(tt.item())

    ms = rmax(ms)
    total_nodes = rsum(float(n_nodes))
    for name in e2e:
        e2e[name][0] = rmax(e2e[name][0])
    launches = int(rsum(float(launches)))
    per_step_ms = {k: rmax(per_step_ms[k]) for k in sorted(per_step_ms)}  # slowest rank per kernel class

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
        f_vis = n_visible / max(n_entries, 1)
        lvl_bytes_per_step = FUSED_BYTES_PER_NODE * n_nodes  # the 4*f_vis list write belongs to k_compact
        achieved = lvl_bytes_per_step * K / (lvl_ms * 1e-3) / 1e9 if lvl_ms > 0 else None
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("k_update_level_bytes_per_launch_sum_per_step")
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": total_nodes * K / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "nodes_per_gpu": n_nodes, "display_entries_per_gpu": n_entries, "visible_fraction": f_vis,
                       "l2": "inputs (%.2f GB per step) larger than the 126 MB L2" % (lvl_bytes_per_step / 1e9), "frame_graph": True,
                       "exchange": None if world == 1 else ("per step, inside the frame graph: k_compact stores the survivors into rank 0's memory over NVLink "
                                   "(peer stores), count + step flag published on the device, rank 0 concatenates and acknowledges; no host sync "
                                   "(%d entries gathered in the last step)" % gathered[0]) if peer else
                                   ("per step: visible counts all-gathered (NCCL), visible lists gathered to rank 0 "
                                   "(grouped send/recv sized from the counts; %d entries gathered in the last step)" % gathered[0]),
                       "scale": args.scale},
            "e2e": {"value": total_nodes * K / e2e["e2e"][0] / 1e6, "unit": UNIT, "h2d_bytes_per_step": e2e["e2e"][1], "d2h_bytes_per_step": e2e["e2e"][2],
                    "what": "every step: hgb_upload_trs of ALL nodes from pinned host + hgb_update + hgb_get_visible to pinned host"},
            "e2e_resident": {"value": total_nodes * K / e2e["e2e_resident"][0] / 1e6, "unit": UNIT, "h2d_bytes_per_step": e2e["e2e_resident"][1],
                             "d2h_bytes_per_step": e2e["e2e_resident"][2], "what": "scene resident (Scene::Update(dt) drop-in): clocks up, visible list down"},
            "gpu_launches": launches,
            "roofline": {"kernel": "k_update_level<fused>", "bound": "hbm", "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": achieved / peak if achieved else None, "traffic": traffic, "algorithmic_bytes_per_node": FUSED_BYTES_PER_NODE,
                         "launches_per_step": lvl_n / K, "kernel_ms_per_step": per_step_ms},
            "clocks": clocks_rec,
        }
        if not args.no_cpu_baseline and world == 1:
            rate, kind, sample = cpu_reference_rate(1, 1, 1_000_000, 40)  # ~10 s of one host core
            line["cpu_baseline"] = {"value": rate / 1e6, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample}
        print(json.dumps(line), flush=True)
    barrier()  # no rank frees its inbox / mailbox while a peer may still write to it
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
