"""6-view frame of the c4 batch (view + 4 PSSM-like splits + spot, forward_pipeline.cpp:331-332,380-381): per-kernel times with all views
compacted in one pass (default) or one k_compact launch per view (HGB_MULTI_COMPACT=0), and the 1-view frame beside them."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import harfang3d_b200
from harfang3d_b200 import hostmirror, scenegen

fs = scenegen.config_c4(seed=4, n_scenes=16, nodes_per_scene=1_000_000, animated_frac=0.25, n_keys=16)
ctx = harfang3d_b200.Context(0)
hostmirror.upload(ctx, fs)
clocks = hostmirror.PlayClocks(fs)
cam = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
views6 = np.concatenate([hostmirror.scene_frusta(fs)[0:1]] + [hostmirror.make_perspective_frustum(n, f, 1.0, 16.0, 9.0, cam) for n, f in
                        ((0.1, 10.0), (10.0, 50.0), (50.0, 100.0), (100.0, 200.0))] + [hostmirror.make_perspective_frustum(0.1, 80.0, 0.6, 1.0, 1.0, cam)])
out = {}
for name, v in (("views_1", views6[0:1]), ("views_6", views6)):
    ctx.set_views(np.ascontiguousarray(v, np.float32))
    for _ in range(5):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    ctx.sync()
    ctx.time_begin()
    for _ in range(50):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    ms = ctx.time_end() / 50
    ctx.set_profiling(True); ctx.get_profile(reset=True)
    for _ in range(20):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    prof = ctx.get_profile(reset=True); ctx.set_profiling(False)
    out[name] = {"frame_ms": ms, "k_compact_ms": prof["k_compact"][0] / 20, "k_compact_launches": prof["k_compact"][1] / 20,
                 "k_update_level_ms": prof["k_update_level"][0] / 20, "visible": [int(ctx.get_visible_count(0, i)) for i in range(len(v))]}
out["multi_compact"] = os.environ.get("HGB_MULTI_COMPACT", "1")
print(json.dumps(out))
