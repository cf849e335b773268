#!/usr/bin/env python
"""Roofline of the SECONDARY kernels of the path, each timed alone with the per-kernel CUDA-event counters of the C ABI
(hgb_set_profiling) on the full-size workloads: the stand-alone cull predicate, the order-preserving compaction, the scene
bounding-box reduction, the sort keys (c4: 16M nodes) and the track sampler + bone matrices of c3 (4096 characters x 64 bones).
Algorithmic bytes per unit follow DESIGN.md §5 / SURVEY.md §8(d). Prints one JSON object; run on the GPU box (scripts/gpu_kernels.sh)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import harfang3d_b200  # noqa: E402
from harfang3d_b200 import hostmirror, scenegen  # noqa: E402

REPS = int(os.environ.get("REPS", "20"))
SCALE = float(os.environ.get("SCALE", "1.0"))


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def timed(ctx, fn, cls, reps=REPS):
    """average ms per launch of kernel class `cls` over reps calls of fn (L2 flushed before each call)"""
    ctx.set_profiling(True)
    ctx.get_profile(reset=True)
    for _ in range(reps):
        ctx.flush_l2()
        fn()
    prof = ctx.get_profile(reset=True)
    ctx.set_profiling(False)
    ms, n = prof[cls]
    return ms / max(n, 1), n // reps


def main():
    pk, pk_src = peak()
    out = {"peak_gbs": pk, "peak_source": pk_src, "reps": REPS, "kernels": {}}

    def add(name, units, bytes_per_unit, ms, launches, what):
        gbs = units * bytes_per_unit / (ms * 1e-3) / 1e9
        out["kernels"][name] = {"units": int(units), "algorithmic_bytes_per_unit": bytes_per_unit, "ms_per_launch": ms, "launches_per_call": int(launches),
                                "achieved_gbs": gbs, "frac": gbs / pk, "what": what}

    # ---- c4: 16 scenes x 1M nodes -------------------------------------------------------------------------------------------------
    n_scenes = max(1, int(round(16 * SCALE)))
    fs = scenegen.config_c4(seed=4, n_scenes=n_scenes, nodes_per_scene=1_000_000, animated_frac=0.25, n_keys=16)
    ctx = harfang3d_b200.Context(0)
    t = hostmirror.upload(ctx, fs)
    clocks = hostmirror.PlayClocks(fs)
    for _ in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    n_ent = len(t["ent0_mtx_slot"])
    f_vis = ctx.get_visible_count(0, 0) / max(n_ent, 1)
    ms, nl = timed(ctx, ctx.cull, "k_cull_flags")
    add("k_cull_flags", n_ent, 80, ms, nl, "stand-alone CullModelDisplayLists predicate: 48 B world + 24 B bounds + 8 B entry read per entry (vis byte not counted)")
    ms, nl = timed(ctx, ctx.cull, "k_compact")
    add("k_compact", n_ent, 1 + 4 * f_vis, ms, nl, "order-preserving compaction: 1 B flag read + 4 B per survivor written (f_vis = %.3f)" % f_vis)
    slots, ob_idx, ob, seg, n_seg = hostmirror.object_table(fs)
    ctx.set_objects(slots, ob_idx, ob, seg, n_seg)
    ctx.minmax()
    ms, nl = timed(ctx, lambda: ctx.minmax(fetch=False), "k_minmax")
    add("k_minmax", len(slots), 56, ms, nl, "Scene::GetMinMax reduction: 48 B world + 4 B position + 4 B model per object node")
    view = np.ascontiguousarray(hostmirror.inverse_fast(np.asarray(fs.camera["world12"], np.float32).reshape(1, 12))[0])
    n_vis = ctx.get_visible_count(0, 0)
    ms, nl = timed(ctx, lambda: ctx._ck(ctx._L.hgb_sort_keys(ctx._h, 0, 0, view.ctypes.data)), "k_sort_keys")
    add("k_sort_keys", n_vis, 88, ms, nl, "ComputeModelDisplayListSortKeys of the survivors: 4 B index + 8 B entry + 48 B world + 24 B bounds read, 4 B key written (gather)")
    ctx.close()
    del fs, t

    # ---- c3: 4096 characters x 64 bones, 3 Hermite tracks per bone (half the characters: quaternion rotation) ----------------------------
    n_chars = max(8, int(round(4096 * SCALE)))
    fs = scenegen.config_c3(seed=3, n_chars=n_chars, n_bones=64, n_keys=32)
    ctx = harfang3d_b200.Context(0)
    t = hostmirror.upload(ctx, fs)
    clocks = hostmirror.PlayClocks(fs)
    n_tracks = len(fs.trk_anim)

    def frame():
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
        ctx.bone_matrices(with_previous=True)
    for _ in range(3):
        frame()
    ms, nl = timed(ctx, frame, "k_sample_tracks")
    add("k_sample_tracks(c3)", n_tracks, 172, ms * nl, 1, "3 tracks per bone (rounds = launches per frame: %d), K = 32 keys: SURVEY figure 5*8 probes + 4*32 keys + 12 written ~ 180 B; 172 B used as in DESIGN.md §5" % nl)
    ms, nl = timed(ctx, frame, "k_bone_matrices")
    add("k_bone_matrices(c3)", ctx.n_bones, 336, ms, nl, "bone matrices with the previous-frame twin: 2 x 48 B world + 48 B bind pose + 8 B indices read, 2 x 64 B written")
    ms, nl = timed(ctx, frame, "k_update_level")
    add("k_update_level(c3)", fs.n, 120, ms * nl, 1, "fused level pass of the character batch, all %d levels of a frame (all nodes animated: + 3 x 16 B sampled triples not counted)" % nl)
    ctx.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
