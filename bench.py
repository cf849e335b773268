#!/usr/bin/env python
"""bench.py — scene nodes updated+culled per second on B200 (BASELINE.json metric), one JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c1|c2|c3|c5] [--strong] [--impl reference]

A "step" is one frame of the hot path over one batch of synthetic scenes: Scene::Update (SceneAnim key sampling + world-matrix
propagation) fused with GetModelDisplayLists/CullModelDisplayLists (visibility + order-preserving compaction); c3 adds the skin
bone matrices with their previous-frame twin.
Default workload at every N (weak scaling, per-GPU work fixed): BASELINE.json configs[3] — the 16M-node mixed static/animated scene
batch (16 scenes x 1M nodes, depth 8, 25 % of the nodes driven by a Hermite Position track) — the configuration the metric's roofline
target is quoted on; every rank owns its own batch of independent scene instances (different seed), and at N>1 the visible lists are
gathered to rank 0 inside every step (the path has no other exchange). The other BASELINE.json configurations are --workload c1
(configs[0]), c2 (configs[1]), c3 (configs[2]) and c5 [--strong] (configs[4]: 8192 instances partitioned over the ranks).

  value          nodes/s with the scene resident in HBM, CUDA events on the launching stream, max over ranks
  e2e            the same metric through the C ABI (the drop-in boundary, include/hgb200.h) with HOST buffers inside the timed region,
                 scene resident like the reference's hg::Scene: every step hgb_set_views (frusta) + hgb_update (play clocks) from host
                 memory, hgb_get_visible into pinned host memory (the visible index list)
  e2e_facade     one layer up, the call a HARFANG user makes: hg::Scene::Update(dt) + GetVisibleModelDisplayLists through the C++ facade,
                 result = the std::vector<ModelDisplayList> the reference arm's GetModelDisplayLists + CullModelDisplayLists leave behind,
                 materialised on the host (16 bytes per visible entry gathered by host threads)
  e2e_full_upload  as e2e, plus hgb_upload_trs of EVERY node from pinned host memory each step (all transforms host-edited: the
                 physics / script-driven worst case; PCIe-bound)
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
# algorithmic bytes (SURVEY.md §8d / DESIGN.md §5)
FUSED_BYTES_PER_NODE = 120       # fused update+cull: TRS 36 + parent 4 + entry 8 + AABB 24 read, world 48 written (+ 4 B per survivor: k_compact)
BONE_BYTES = 336                 # world 48 + bind pose 48 + indices 8 read, 64 written; x2 with the previous-frame twin


def track_bytes(n_keys):
    return int(np.ceil(np.log2(max(n_keys, 2)))) * 8 + 4 * 32 + 12  # time probes + 4 keys + TRS lane write


WORKLOAD_TEXT = {
    "c4": "c4: %d scenes x 1M nodes, depth 8, 25%% animated (Position track, 16 keys), fused update+cull, 1 camera",
    "c2": "c2: %d-node static scene, depth 8, fused update+cull",
    "c5": "c5: %d instances x 2048 nodes per GPU, depth 6, one camera per instance",
    "c5s": "c5 strong: %d instances x 2048 nodes split over %d GPU(s), depth 6, one camera per instance, visibility gather to rank 0",
    "c3": "c3: %d skinned characters x 64 bones (+ object and prop nodes), Position/Rotation/Scale tracks (32 keys, half quaternion), "
          "key sampling + world matrices + cull + bone matrices with previous-frame twin",
    "c1": "c1: 10k-node static scene, 4 levels, fused update+cull",
}


def workload_text(name, world, scale, strong):
    """the config.workload string — the same for both arms (the driver compares them)"""
    if name == "c4":
        return WORKLOAD_TEXT["c4"] % max(1, int(round(16 * scale)))
    if name == "c2":
        return WORKLOAD_TEXT["c2"] % max(1000, int(1_000_000 * scale))
    if name == "c1":
        return WORKLOAD_TEXT["c1"]
    if name == "c3":
        return WORKLOAD_TEXT["c3"] % max(1, int(4096 * scale))
    if strong and world > 1:
        return WORKLOAD_TEXT["c5s"] % (max(world, int(8192 * scale)), world)
    return WORKLOAD_TEXT["c5"] % max(1, int(4096 * scale))


def workload(name, rank, world=1, scale=1.0, strong=False):
    """-> (FlatScene, keys per track). scale < 1 shrinks the node count (smoke runs / CPU samples)."""
    if name == "c4":
        n_scenes = max(1, int(round(16 * scale)))
        return scenegen.config_c4(seed=4 + 1000 * rank, n_scenes=n_scenes, nodes_per_scene=1_000_000, animated_frac=0.25, n_keys=16), 16
    if name == "c2":
        return scenegen.config_c2(seed=2 + 1000 * rank, n_nodes=max(1000, int(1_000_000 * scale))), 0
    if name == "c1":
        return scenegen.config_c1(seed=1 + 1000 * rank), 0
    if name == "c3":
        return scenegen.config_c3(seed=3 + 1000 * rank, n_chars=max(1, int(4096 * scale))), 32
    if name == "c5" and strong and world > 1:
        # BASELINE configs[4] as stated: ONE batch of 8192 instances, contiguous blocks per rank (sharding.partition_instances)
        from harfang3d_b200 import sharding
        total = max(world, int(8192 * scale))
        full = scenegen.config_c5(seed=5, n_instances=total)
        b = sharding.partition_instances(total, world)
        return scenegen.slice_instances(full, int(b[rank]), int(b[rank + 1])), 0
    if name == "c5":
        # per-rank share of BASELINE configs[4] at 2 GPUs (8192 instances x 2048 nodes = 2^24 slots exceed the reference's 2^24-1
        # vector_list capacity for ONE scene/context, foundation/vector_list.h:40)
        return scenegen.config_c5(seed=5 + 1000 * rank, n_instances=max(1, int(4096 * scale))), 0
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
def cpu_sample_scene(name, i, seed, nodes_per_scene):
    """scene i of the CPU sample of a workload: c4 = one of its 1M-node scenes; the small / skinned workloads = the workload itself"""
    if name == "c4":
        return scenegen.config_c4(seed=seed + i, n_scenes=1, nodes_per_scene=nodes_per_scene, animated_frac=0.25, n_keys=16)
    if name == "c2":
        return scenegen.config_c2(seed=2 + i, n_nodes=nodes_per_scene)
    if name == "c1":
        return scenegen.config_c1(seed=1 + i)
    if name == "c3":
        return scenegen.config_c3(seed=3 + i, n_chars=max(1, nodes_per_scene // 66))
    if name == "c5":  # instances are independent hg::Scenes in the reference: a sample scene = ONE instance with its own camera
        fs = scenegen.config_c5(seed=5 + i, n_instances=1)
        return fs
    raise SystemExit("no CPU sample for workload %s" % name)


def cpu_reference_rate(threads, n_scenes, nodes_per_scene, frames, seed=4, warm=1, name="c4"):
    """nodes/s of the reference's own Scene::Update + GetModelDisplayLists + CullModelDisplayLists (+ bone matrices for c3) on `threads`
    host threads. The sample is `n_scenes` scenes of the workload (one hg::Scene each — the reference path itself is single-threaded,
    scene instances are independent), dealt round-robin to the threads; a frame = every scene stepped once.
    Returns (rate, kind, description, phases): phases = seconds per frame in {Update, GetModelDisplayLists, Cull, bones} summed over the
    sample's scenes (reference build only)."""
    from tests import oracle_lib
    from tests.oracle_lib import P
    ref = oracle_lib.load_ref()
    kind = "reference" if ref is not None else "port"
    threads = max(1, min(threads, n_scenes))
    scenes = [None] * n_scenes
    phase_acc = [np.zeros(4) for _ in range(threads)]
    with_bones = 1 if name == "c3" else 0
    nodes_of = [0] * n_scenes

    def build(th):
        for i in range(th, n_scenes, threads):
            fs = cpu_sample_scene(name, i, seed, nodes_per_scene)
            nodes_of[i] = fs.n
            fr = np.ascontiguousarray(np.asarray(hostmirror.scene_frusta(fs), np.float32).reshape(-1, 24)[0])
            if ref is not None:
                from tests.scene_io import RefScene
                scenes[i] = (RefScene(ref, fs), fr)
            else:
                from tests.port_scene import PortScene
                scenes[i] = (PortScene(oracle_lib.load_port(), fs), fr)

    def run(th, n_frames, record):
        for _ in range(n_frames if ref is None else 1):
            for i in range(th, n_scenes, threads):
                sc, fr = scenes[i]
                if ref is not None:  # scenes are independent: n_frames of one scene, then the next, is the same work as frame-major order
                    out = np.zeros(4)
                    ref.hgref_time_frames(sc.h, n_frames, FRAME_DT, P(fr), with_bones, P(out))
                    if record:
                        phase_acc[th] += out
                else:
                    sc.update(FRAME_DT)
                    sc.cull(0, fr)

    def par(fn, *a):
        th = [threading.Thread(target=fn, args=(i,) + a) for i in range(threads)]
        [t.start() for t in th]
        [t.join() for t in th]

    par(build)
    par(run, max(warm, 1), False)  # warm-up frames (the first Update sizes the vectors)
    t0 = time.perf_counter()
    par(run, frames, True)
    wall = time.perf_counter() - t0
    for sc, _ in scenes:
        sc.close()
    total_nodes = int(sum(nodes_of))
    rate = total_nodes * frames / wall
    phases = None
    if ref is not None:
        s = sum(phase_acc) / frames
        phases = {"update_s": float(s[0]), "get_model_display_lists_s": float(s[1]), "cull_s": float(s[2]), "bones_s": float(s[3]),
                  "what": "seconds per frame summed over the sample's scenes (thread-seconds)"}
    return rate, kind, "%d hg::Scene(s), %d nodes in all (scenes of the %s workload) on %d thread(s) x %d frames, wall %.2f s" % (
        n_scenes, total_nodes, name, threads, frames, wall), phases


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_sample(name, scale):
    """(n_scenes, nodes_per_scene) of the CPU sample used by the reference arm"""
    if name == "c4":
        return max(1, int(round(16 * scale))), 1_000_000
    if name == "c2":
        return 16, max(1000, int(1_000_000 * scale))
    if name == "c1":
        return 16, 10_000
    if name == "c3":
        return 16, max(66, int(4096 * 66 * scale) // 16)
    return 256, 2048  # c5: 256 instances, one hg::Scene each


def run_reference_arm(args, rank, world):
    """The reference's own CPU implementation on all host threads, on the workload itself (c4: 16 hg::Scenes of 1M nodes, scaled by
    --scale), one step = one frame of every scene."""
    if rank != 0:
        return
    threads = host_threads()
    n_scenes, nodes = reference_sample(args.workload, args.scale)
    rate, kind, sample, phases = cpu_reference_rate(threads, n_scenes, nodes, args.steps, warm=args.warmup, name=args.workload)
    v = rate / 1e6
    fs_nodes = {"c4": n_scenes * nodes, "c2": max(1000, int(1_000_000 * args.scale)), "c1": 10_000, "c3": max(1, int(4096 * args.scale)) * 66}.get(
        args.workload, 2048 * (max(world, int(8192 * args.scale)) if (args.strong and world > 1) else max(1, int(4096 * args.scale))))
    cb = {"value": v, "unit": UNIT, "cores": min(threads, n_scenes), "kind": kind, "sample": sample}
    if phases:
        cb["phases"] = phases
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": fs_nodes / rate * 1e3,  # one frame of the whole workload at the measured rate
            "higher_is_better": True, "scaling": "strong" if (args.strong and world > 1) else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": workload_text(args.workload, world, args.scale, args.strong), "scale": args.scale},
            "cpu_baseline": cb,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def list_checksum(idx):
    """order-sensitive 63-bit checksum of an index list"""
    a = np.asarray(idx, np.uint64)
    w = np.arange(1, len(a) + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return int((a * w + (a ^ (w << np.uint64(17)))).sum(dtype=np.uint64)) & 0x7fffffffffffffff


# ---- own arm ------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default="c4", choices=["c4", "c1", "c2", "c3", "c5"])
    ap.add_argument("--strong", action="store_true", help="c5 only: ONE batch of 8192 instances partitioned over the ranks (BASELINE configs[4]) instead of 4096 per rank")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; the reported config says so)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-facade", action="store_true", help="skip the hg::Scene facade leg (e2e then is the C-ABI leg)")
    ap.add_argument("--lookahead", type=int, default=0, help="A/B switch: L2 prefetch distance in positions (0 = default, -1 = off)")
    ap.add_argument("--tma-level", action="store_true", help="A/B switch: persistent TMA-staged level kernel instead of one thread per node (LDG rows)")
    ap.add_argument("--top-walk", type=int, default=0, help="A/B switch: budget (node evaluations) of the one-launch top of the hierarchy (0 = default, -1 = one launch per level)")
    ap.add_argument("--slice-nodes", type=int, default=0, help="A/B switch: slice size of big levels under the overlapped sampler (0 = no slicing)")
    ap.add_argument("--overlap", action="store_true", help="A/B switch: sample tracks by target level on a second stream, overlapped with the level kernels")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"], help="N>1 result gather: peer = survivors stored into rank 0's memory over NVLink "
                    "by the compaction kernel itself, flags + acks on the device, no host sync (hgb_peer_*); nccl = counts all-gathered, host-sized send/recv (A/B)")
    ap.add_argument("--async-post", type=int, default=None, choices=[0, 1], help="A/B switch: compaction + gather of a frame as its own graph on a second stream, "
                    "overlapping the next frame (hgb_options.async_post; default on)")
    ap.add_argument("--row-store", action="store_true", help="A/B switch: per-thread world row stores instead of the transposed warp store")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.strong and args.workload != "c5":
        raise SystemExit("--strong applies to --workload c5")

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import harfang3d_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this path has no CPU fallback (use --impl reference for the CPU reference)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    fs, n_keys = workload(args.workload, rank, world, args.scale, args.strong)
    desc = workload_text(args.workload, world, args.scale, args.strong)
    n_nodes = fs.n
    is_c3 = args.workload == "c3"
    ctx = harfang3d_b200.Context(local_rank, row_store=args.row_store, lookahead=args.lookahead, tma_level=args.tma_level, top_walk=args.top_walk,
                                  overlap_sampler=args.overlap, slice_nodes=args.slice_nodes,
                                  # the NCCL A/B reads the lists on the caller's stream: no asynchronous tail there
                                  async_post=(False if (world > 1 and args.exchange == "nccl") else None) if args.async_post is None else bool(args.async_post))
    t = hostmirror.upload(ctx, fs)
    n_entries = len(t["ent0_mtx_slot"]) + len(t["ent1_mtx_slot"])
    n_bones = len(t["skin_mtx_slot"])
    n_tracks = len(t["tracks"])
    clocks = hostmirror.PlayClocks(fs)
    K, W = args.steps, max(args.warmup, 3)
    frusta = np.ascontiguousarray(hostmirror.scene_frusta(fs), np.float32)

    # N>1: the path's only exchange — visible lists gathered to rank 0
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
        inst_all = torch.zeros(world, dtype=torch.int64, device="cuda")
        inst_all[rank] = len(fs.instance_cameras) if fs.instance_cameras is not None else 0
        dist.all_reduce(inst_all)
        inst_all = inst_all.cpu().numpy()
    if peer:
        # the exchange rides on the frame itself: k_compact stores the survivors into rank 0's inbox over NVLink, flags and
        # acknowledgements are device words, rank 0 concatenates — nothing of it touches the host (hgb_kernels_peer.cuh)
        aux = inst_all.astype(np.uint32) if fs.instance_cameras is not None else None
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
            bounds = np.concatenate([[0], np.cumsum(inst_all)]).astype(np.int64)
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

    # the play clocks of the steps are stepped ahead of the timed loops when there are many (exact int64 host arithmetic,
    # hostmirror.PlayClocks): a Python loop over thousands of playing anims (c3) must not sit between two frames of a device-timed region
    clock_ring = [clocks.step(FRAME_DT) for _ in range(64)] if len(fs.plays) > 64 else None
    step_no = [0]

    def next_clocks():
        step_no[0] += 1
        return clock_ring[step_no[0] % 64] if clock_ring is not None else clocks.step(FRAME_DT)

    def step_resident():
        ctx.update(next_clocks(), fused_cull=True)
        if is_c3:
            ctx.bone_matrices(with_previous=True)
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

    # ---- N>1: parity word of the exchange (outside the timed region): what rank 0 holds == what every rank computed ------------
    exchange_check = None
    if world > 1 and peer:
        step_p, err = ctx.peer_status()
        if err:
            raise SystemExit("bench.py: peer gather error word %d on rank %d after %d frames" % (err, rank, step_p))
        mine = ctx.get_visible(0, 0)
        rec = torch.tensor([len(mine), list_checksum(mine.astype(np.uint64) + np.uint64(entry_base[rank]))], dtype=torch.int64, device="cuda")
        recs = [torch.zeros(2, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(recs, rec)
        ok = True
        if rank == 0:
            off, idx, _aux = ctx.peer_result()
            gathered[0] = int(off[-1])
            ok = len(idx) == int(off[-1])
            for r in range(world):
                cnt, chk = int(recs[r][0].item()), int(recs[r][1].item())
                seg = idx[int(off[r]):int(off[r + 1])]
                ok = ok and len(seg) == cnt and list_checksum(seg) == chk
            exchange_check = {"ok": bool(ok), "entries": int(off[-1]), "ranks": world,
                              "what": "per-rank visible count + order-sensitive checksum of the local list (globalised) == the same over rank r's segment of the gathered list"}
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.broadcast(flag, 0)
        if not int(flag.item()):
            if rank == 0:
                print(json.dumps({"exchange_check": exchange_check}), file=sys.stderr)
            raise SystemExit("bench.py: exchange_check failed — the gathered list differs from the per-rank lists")

    # ---- per-kernel CUDA-event times over more steps (plain launches, event pair around every kernel) ---------------------
    KP = min(K, 50)
    ctx.set_profiling(True)
    ctx.get_profile(reset=True)
    for _ in range(KP):
        step_resident()
    prof = ctx.get_profile(reset=True)
    ctx.set_profiling(False)
    lvl_ms, lvl_n = prof["k_update_level"]
    per_step_ms = {k: v[0] / KP for k, v in prof.items() if v[1]}

    # ---- end to end with host buffers -------------------------------------------------------------------------------------
    pin_out = torch.empty(max(len(t["ent0_mtx_slot"]), 1), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    e2e = {}
    KE = min(K, 50)
    legs = [("e2e_capi", False)]
    if args.workload in ("c4", "c2") and world == 1:
        legs.append(("e2e_full_upload", True))
    for name, full in legs:
        pin_trs = None
        if full:
            pin_trs = torch.empty((n_nodes, 9), dtype=torch.float32, pin_memory=True).numpy()
            pin_trs[:] = fs.trs

        def step_capi():
            if pin_trs is not None:
                ctx.upload_trs(pin_trs)
            if fs.instance_cameras is None:
                ctx.set_views(frusta)  # the camera of this frame (96 bytes per view)
            step_resident()
            return ctx.get_visible(0, 0, out=pin_out)

        for _ in range(6):  # (the library settles on its frame mode for this calling pattern within three steps; both parities captured by the fifth)
            vis = step_capi()
        barrier()
        t0 = time.perf_counter()
        for _ in range(KE):
            vis = step_capi()
        ctx.sync()
        dt_s = time.perf_counter() - t0
        barrier()
        h2d = (n_nodes * 36 if full else 0) + 8 * len(fs.plays) + (0 if fs.instance_cameras is not None else frusta.size * 4)
        e2e[name] = [dt_s / KE, h2d, 4 + 4 * len(vis), None]
        del pin_trs

    clocks_rec = sampler.stop()  # sampled from the start of the timed region to the end of the C-ABI e2e legs (all under load)

    # ---- e2e through the hg::Scene facade: Scene::Update(dt) + GetVisibleModelDisplayLists (the drop-in a HARFANG user calls) ------
    facade_note = None
    if not args.no_facade and fs.instance_cameras is None and not is_c3:
        try:
            from harfang3d_b200 import flatscene  # ctypes table of the product's flat binding over the facade
            fac = flatscene.load()
            fac.hgfacade_set_default_device(local_rank)  # one process per GPU: the facade's scenes live on this rank's device
            sc = flatscene.SceneHandle(fac, fs)
            fr0 = np.ascontiguousarray(frusta[0])
            out3 = np.zeros(3)
            fac.hgfacade_time_fused_frames(sc.h, 5, FRAME_DT, flatscene.P(fr0), flatscene.P(out3))  # warm-up (the first frame uploads the scene)
            # no collective inside this leg (a rank that fails it must not strand the others): every call below ends with the list on the host
            bd = np.zeros(3)
            fac.hgfacade_visible_breakdown(flatscene.P(bd))  # clear what the warm-up accumulated
            nvis = fac.hgfacade_time_fused_frames(sc.h, KE, FRAME_DT, flatscene.P(fr0), flatscene.P(out3))
            fac.hgfacade_visible_breakdown(flatscene.P(bd))
            e2e["e2e_facade"] = [float(out3[2]) / KE, 8 * len(fs.plays) + 96, 4 + 4 * int(nvis),
                                 {"update_ms": float(out3[0]) / KE * 1e3, "visible_lists_ms": float(out3[1]) / KE * 1e3,
                                  "visible_lists_breakdown_ms": {"device_round_trip": float(bd[0]) / KE * 1e3, "vector_resize": float(bd[1]) / KE * 1e3,
                                                                 "copy_into_vector": float(bd[2]) / KE * 1e3}, "visible": int(nvis),
                                  "list_bytes_materialised_on_host": 16 * int(nvis)}]
            sc.close()
        except Exception as e:  # the line says why there is no facade number
            facade_note = "%s: %s" % (type(e).__name__, e)
    torch.cuda.set_device(local_rank)
    if args.no_facade:
        pass
    elif facade_note is None and "e2e_facade" not in e2e:
        facade_note = "the facade leg covers the single-camera rigid workloads (c4, c2, c1); %s is measured through the C ABI" % args.workload

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
    have_facade = rsum(1.0 if "e2e_facade" in e2e else 0.0) == world
    if not have_facade:
        e2e.pop("e2e_facade", None)
    for name in sorted(e2e):
        e2e[name][0] = rmax(e2e[name][0])
    launches = int(rsum(float(launches)))
    per_step_ms = {k: rmax(per_step_ms.get(k, 0.0)) for k in sorted(per_step_ms)}  # slowest rank per kernel class
    lvl_ms = rmax(lvl_ms)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
        f_vis = n_visible / max(n_entries, 1)

        def leg(name, what):
            s_per_step, h2d, d2h, extra = e2e[name]
            o = {"value": total_nodes / s_per_step / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": s_per_step * 1e3, "what": what}
            if extra:
                o.update(extra)
            return o

        what_capi = "C ABI, scene resident: hgb_set_views (frusta) + hgb_update (play clocks) from host memory, hgb_get_visible into pinned host memory (index list)"
        what_fac = ("hg::Scene facade, scene resident like the reference's hg::Scene: Scene::Update(dt) + GetVisibleModelDisplayLists -> std::vector<ModelDisplayList> "
                    "filled on the host (what GetModelDisplayLists + CullModelDisplayLists leave behind in the reference arm)")
        lvl_bytes_per_step = FUSED_BYTES_PER_NODE * n_nodes  # the 4*f_vis list write belongs to k_compact
        achieved = lvl_bytes_per_step * KP / (lvl_ms * 1e-3) / 1e9 if lvl_ms > 0 else None
        traffic, traffic_src = None, None
        if args.workload == "c4" and args.scale == 1.0:
            try:
                tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
                traffic = tj.get("k_update_level_bytes_per_launch_sum_per_step")
                traffic_src = "profiles/traffic.json (%s; ncu --set full capture of the same command, not this run)" % tj.get("capture", "ncu capture")
            except Exception:
                pass
        roof = {"kernel": "k_update_level<fused> (+ k_update_top: the leading levels in one launch)", "bound": "hbm", "achieved": achieved, "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak if achieved else None, "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes_per_node": FUSED_BYTES_PER_NODE, "launches_per_step": lvl_n / KP, "kernel_ms_per_step": per_step_ms,
                "timing": "CUDA-event pair around every launch, plain (unchained) launches, %d steps; ms_per_step of the line is the chained frame graph, "
                          "whose kernels overlap their ramps: the per-kernel times sum to more than it" % KP}
        others = []
        if n_tracks and per_step_ms.get("k_sample_tracks"):
            a = track_bytes(n_keys) * n_tracks / (per_step_ms["k_sample_tracks"] * 1e-3) / 1e9
            others.append({"kernel": "k_sample_tracks", "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                           "algorithmic_bytes_per_track": track_bytes(n_keys), "tracks": n_tracks})
        if n_bones and per_step_ms.get("k_bone_matrices"):
            a = BONE_BYTES * n_bones / (per_step_ms["k_bone_matrices"] * 1e-3) / 1e9
            others.append({"kernel": "k_bone_matrices (with previous-frame twin)", "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                           "algorithmic_bytes_per_bone": BONE_BYTES, "bones": n_bones})
        line = {
            "metric": METRIC, "value": total_nodes * K / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if (args.strong and world > 1) else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": desc, "nodes_per_gpu": n_nodes, "display_entries_per_gpu": n_entries, "visible_fraction": f_vis,
                       "l2": ("inputs (%.2f GB per step) larger than the 126 MB L2" % (lvl_bytes_per_step / 1e9)) if lvl_bytes_per_step > 126e6 else
                             ("inputs (%.3f GB per step) fit the 126 MB L2 and no flush is issued between steps: this small configuration is a "
                              "launch/latency-bound parity case, not the roofline line" % (lvl_bytes_per_step / 1e9)),
                       "frame_graph": True,
                       "exchange": None if world == 1 else ("per step, inside the frame: k_compact stores the survivors into rank 0's memory over NVLink "
                                   "(peer stores), count + step flag published on the device, rank 0 concatenates and acknowledges; no host sync "
                                   "(%d entries gathered in the last step)" % gathered[0]) if peer else
                                   ("per step: visible counts all-gathered (NCCL), visible lists gathered to rank 0 "
                                   "(grouped send/recv sized from the counts; %d entries gathered in the last step)" % gathered[0]),
                       "scale": args.scale},
            "e2e": leg("e2e_capi", what_capi),
            "gpu_launches": launches,
            "roofline": roof,
            "clocks": clocks_rec,
        }
        if "e2e_facade" in e2e:
            line["e2e_facade"] = leg("e2e_facade", what_fac)
        if "e2e_full_upload" in e2e:
            line["e2e_full_upload"] = leg("e2e_full_upload", "as e2e plus hgb_upload_trs of ALL nodes from pinned host memory every step (every transform host-edited; PCIe-bound)")
        if facade_note:
            line["e2e_facade_note"] = facade_note
        if others:
            line["roofline_others"] = others
        if exchange_check is not None:
            line["exchange_check"] = exchange_check
        if not args.no_cpu_baseline and world == 1:
            if args.workload == "c4":
                rate, kind, sample, phases = cpu_reference_rate(1, 1, 1_000_000, 40)  # ~10 s of one host core
            elif args.workload == "c1":
                rate, kind, sample, phases = cpu_reference_rate(1, 1, 10_000, 200, name="c1")
            else:
                _, nn = reference_sample(args.workload, args.scale)
                rate, kind, sample, phases = cpu_reference_rate(1, 1, nn, 20 if nn > 200_000 else 200, name=args.workload)
            line["cpu_baseline"] = {"value": rate / 1e6, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample}
            if phases:
                line["cpu_baseline"]["phases"] = phases
        print(json.dumps(line), flush=True)
    barrier()  # no rank frees its inbox / mailbox while a peer may still write to it
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
