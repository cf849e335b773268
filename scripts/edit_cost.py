"""Cost of a hierarchy edit: hgb_set_topology (full depth re-sort: host BFS + device gather/scatter of TRS and both
world buffers) and the first hgb_update after it (tables + graph re-captured), at 1M and 16M nodes, for one re-parented leaf."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import harfang3d_b200
from harfang3d_b200 import hostmirror, scenegen

out = {}
for name, fs in (("1M", scenegen.config_c2(seed=2, n_nodes=1_000_000)), ("16M", scenegen.config_c4(seed=4, n_scenes=16, nodes_per_scene=1_000_000, animated_frac=0.25, n_keys=16))):
    ctx = harfang3d_b200.Context(0)
    t = hostmirror.upload(ctx, fs)
    clocks = hostmirror.PlayClocks(fs)
    for _ in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    ctx.sync()
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); steady = time.perf_counter() - t0
    parent = t["parent_slot"].copy()
    leaf = fs.n - 1
    parent[leaf] = parent[leaf - 5]  # re-parent one leaf inside its level
    t0 = time.perf_counter(); ctx.set_topology(parent); ctx.sync(); topo = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); first = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); second = time.perf_counter() - t0
    # the incremental path: the same leaf moved under another parent of its level (hgb_set_parents: parent rows only)
    slot = np.array([leaf], np.uint32)
    newp = np.array([parent[leaf - 9]], np.uint32)
    t0 = time.perf_counter(); rebuilt = ctx.set_parents(slot, newp); ctx.sync(); inc = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); inc_first = time.perf_counter() - t0
    out[name] = {"incremental_reparent_ms": inc * 1e3, "incremental_rebuilt": bool(rebuilt), "first_update_after_incremental_ms": inc_first * 1e3,"steady_frame_ms": steady * 1e3, "set_topology_ms": topo * 1e3, "first_update_after_edit_ms": first * 1e3, "second_update_ms": second * 1e3}
    ctx.close()
    # in-place spawn / despawn (hgb_options.append_slack): one leaf with a display entry appended under a deep node, then removed again
    ctx = harfang3d_b200.Context(0, append_slack=65536)
    t = hostmirror.upload(ctx, fs)
    clocks = hostmirror.PlayClocks(fs)
    for _ in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    ctx.sync()
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); steady_slack = time.perf_counter() - t0
    slot = np.array([fs.n], np.uint32)
    par = np.array([fs.n - 7], np.uint32)
    trs = np.array([[1, 2, 3, 0.1, 0.2, 0.3, 1, 1, 1]], np.float32)
    t0 = time.perf_counter()
    r1 = ctx.append_nodes(slot, par, trs)
    r2 = ctx.append_display_entries(0, slot, np.array([0], np.uint32), np.array([0], np.uint8))
    ctx.sync(); app_first_call = time.perf_counter() - t0  # allocates the edit scratch
    slot = slot + 1
    t0 = time.perf_counter()
    r1 = ctx.append_nodes(slot, par, trs) or r1
    r2 = ctx.append_display_entries(0, slot, np.array([0], np.uint32), np.array([0], np.uint8)) or r2
    ctx.sync(); app = time.perf_counter() - t0
    batch = np.arange(int(slot[0]) + 1, int(slot[0]) + 1 + 1024, dtype=np.uint32)  # an episode's worth: 1024 leaves under random nodes
    bpar = np.random.default_rng(1).integers(0, fs.n, 1024).astype(np.uint32)
    t0 = time.perf_counter()
    r1 = ctx.append_nodes(batch, bpar, np.tile(trs, (1024, 1))) or r1
    r2 = ctx.append_display_entries(0, batch, np.zeros(1024, np.uint32), np.zeros(1024, np.uint8)) or r2
    ctx.sync(); app_batch = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); app_first = time.perf_counter() - t0
    t0 = time.perf_counter(); r3 = ctx.remove_nodes(slot); ctx.sync(); rem = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True); ctx.sync(); rem_first = time.perf_counter() - t0
    out[name].update({"append_leaf_with_entry_ms": app * 1e3, "append_first_call_ms": app_first_call * 1e3, "append_1024_leaves_with_entries_ms": app_batch * 1e3, "append_rebuilt": bool(r1 or r2), "first_update_after_append_ms": app_first * 1e3,
                      "remove_leaf_ms": rem * 1e3, "remove_rebuilt": bool(r3), "first_update_after_remove_ms": rem_first * 1e3,
                      "steady_frame_with_append_region_ms": steady_slack * 1e3})
    ctx.close()
print(json.dumps(out))
