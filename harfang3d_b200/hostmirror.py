"""Python mirror of the host-side logic that sits between hg::Scene and the C ABI, for tests and bench.py.

The shipping host side is C++ (csrc/facade: the hg::Scene-compatible facade). This module restates the same edit-time logic over a
FlatScene so Python can drive the C ABI directly:
  * display-entry tables in the reference's emission order — Scene::GetModelDisplayLists, harfang/engine/scene.cpp:318-404
  * skin records (bone -> transform slot, fallback to the object's own transform) — scene.cpp:365-394
  * bound tracks from SceneAnim/Anim — BindNodeAnim scene.cpp:2070-2152, EvaluateBoundAnim :2154-2181
  * play clocks — UpdatePlayingAnims scene.cpp:2380-2431 (int64, exact)
  * view frusta — ComputePerspectiveProjectionMatrix projection.cpp:29-37, MakeFrustum/TransformFrustum frustum.cpp:12-28
No device math happens here; per-frame arithmetic is all in libhgb200.so.
"""
import math

import numpy as np

from .capi import ENTRY_DISABLED, NOT_PLAYING, ROOT, TRACK_DTYPE

f32 = np.float32


# ---- frusta (float32, reference operation order; vectorised over cameras) ---------------------------------------------
def inverse_fast(m12):
    """InverseFast, foundation/matrix4.cpp:293-303. m12: (n,12) f32"""
    m = m12.reshape(-1, 3, 4)
    s = f32(1) / (m[:, 0, :3] * m[:, 0, :3] + m[:, 1, :3] * m[:, 1, :3] + m[:, 2, :3] * m[:, 2, :3])  # per column
    o = np.zeros_like(m)
    for r in range(3):
        for c in range(3):
            o[:, r, c] = m[:, c, r] * s[:, r]
    tx, ty, tz = -m[:, 0, 3], -m[:, 1, 3], -m[:, 2, 3]
    for r in range(3):
        o[:, r, 3] = o[:, r, 0] * tx + o[:, r, 1] * ty + o[:, r, 2] * tz
    return o.reshape(-1, 12)


def fov_to_zoom_factor(fov):
    lo, hi = f32(0.1) / f32(180) * f32(3.1415926535), f32(179.9) / f32(180) * f32(3.1415926535)
    cf = min(max(f32(fov), lo), hi)
    return f32(1) / f32(math.tan(float(cf * f32(0.5))))


def make_perspective_frustum(znear, zfar, fov, ar_x, ar_y, world12):
    """(n,24) f32 frusta for (n,12) camera world matrices; NDC {origin_bottom_left=false, homogeneous_depth=false}"""
    world12 = np.ascontiguousarray(world12, f32).reshape(-1, 12)
    n = world12.shape[0]
    znear, zfar, ar_x, ar_y = f32(znear), f32(zfar), f32(ar_x), f32(ar_y)
    zoom = fov_to_zoom_factor(fov)
    qa = zfar / (zfar - znear)
    qb = -qa * znear
    rows = {"X": np.array([zoom / ar_x, 0, 0, 0], f32), "Y": np.array([0, zoom / ar_y, 0, 0], f32), "Z": np.array([0, 0, qa, qb], f32)}
    w = np.array([0, 0, 1, 0], f32)
    local = []
    for k in ("Y", "X", "Z"):  # Top=W+Y Bottom=W-Y Left=W+X Right=W-X Near=W+Z Far=W-Z
        for p in (w + rows[k], w - rows[k]):
            ln = np.sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2], dtype=f32)
            local.append(-(p / ln))
    inv = inverse_fast(world12).reshape(n, 3, 4)
    i44 = np.zeros((n, 4, 4), f32)
    i44[:, :3, :] = inv
    i44[:, 3, 3] = 1
    out = np.empty((n, 6, 4), f32)
    for k, v in enumerate(local):
        for r in range(4):  # iMt[r][c] = I44[c][r]
            out[:, k, r] = v[0] * i44[:, 0, r] + v[1] * i44[:, 1, r] + v[2] * i44[:, 2, r] + v[3] * i44[:, 3, r]
    return out.reshape(n, 24)


def scene_frusta(fs):
    c = fs.camera
    if fs.instance_cameras is not None:
        return make_perspective_frustum(c["znear"], c["zfar"], c["fov"], c["ar_x"], c["ar_y"], fs.instance_cameras)
    return make_perspective_frustum(c["znear"], c["zfar"], c["fov"], c["ar_x"], c["ar_y"], c["world12"])


def plane_margins(t, worlds12, frustum24, kind=0, sample=None):
    """Diagnostic of SURVEY.md §8(d): how close the cull predicate of the enabled entries comes to flipping. For every entry and plane
    TestVisibility (frustum.cpp:59-79) compares d - r (outside) and d + r (clipped) with -2w, d = n.(mn+mx), r = |n|.(mx-mn) over the
    transformed box (minmax.cpp:226-244). Returns (min |d - r + 2w|, min |d + r + 2w|, entries looked at), computed in float64 from the
    float32 world rows — a statistic for logs (a visible list that differs from the reference's with a margin of 1e-3 is a bug, not rounding),
    not part of the result. `sample`: every n-th entry only."""
    f64 = np.float64
    slot, bidx, flags = t["ent%d_mtx_slot" % kind], t["ent%d_bounds_idx" % kind], t["ent%d_flags" % kind]
    sel = np.nonzero(flags == 0)[0][:: (sample or 1)]
    if len(sel) == 0:
        return float("inf"), float("inf"), 0
    W = np.asarray(worlds12, f32).reshape(-1, 3, 4)[slot[sel]].astype(f64)
    b = np.asarray(t["bounds"], f32)[bidx[sel]].astype(f64)
    lo, hi = W[:, :, :3] * b[:, None, 0:3], W[:, :, :3] * b[:, None, 3:6]
    mn = W[:, :, 3] + np.minimum(lo, hi).sum(2)
    mx = W[:, :, 3] + np.maximum(lo, hi).sum(2)
    fr = np.asarray(frustum24, f32).reshape(6, 4).astype(f64)
    d = (mn + mx) @ fr[:, :3].T
    r = (mx - mn) @ np.abs(fr[:, :3]).T
    return float(np.abs(d - r + 2 * fr[:, 3]).min()), float(np.abs(d + r + 2 * fr[:, 3]).min()), int(len(sel))


# ---- bounding boxes -------------------------------------------------------------------------------------------------------------
def object_table(fs):
    """Arguments of Context.set_objects for Scene::GetMinMax / GetNodesMinMax: every node with an Object whose model has bounds takes
    part (disabled nodes too: GetNodesWithComponent does not look at flags, scene.cpp:604-640); one MinMax per model = the Union of its
    lists' bounds (Scene::GetObjectMinMax, component.cpp:778-797); one segment per scene instance."""
    nodes = np.nonzero(fs.model >= 0)[0]
    ob = np.zeros((max(len(fs.models), 1), 6), f32)
    has = np.zeros(max(len(fs.models), 1), bool)
    for mi, m in enumerate(fs.models):
        b = np.asarray(m["bounds"], f32).reshape(-1, 6)
        if len(b):
            ob[mi, :3], ob[mi, 3:], has[mi] = b[:, :3].min(0), b[:, 3:].max(0), True
    nodes = nodes[has[fs.model[nodes]]]
    seg, n_seg = None, 1
    if fs.instance_start is not None:
        seg = (np.searchsorted(fs.instance_start, nodes, side="right") - 1).astype(np.uint32)
        n_seg = len(fs.instance_start) - 1
    return nodes.astype(np.uint32), fs.model[nodes].astype(np.uint32), ob, seg, n_seg


# ---- tables ---------------------------------------------------------------------------------------------------------------
def build_tables(fs):
    """FlatScene -> dict of C-ABI arrays. Node i owns transform slot i."""
    n = fs.n
    t = {}
    t["parent_slot"] = np.where(fs.parent >= 0, fs.parent, ROOT).astype(np.uint32)
    # bounds table: one MinMax per (model, list)
    bounds_base, acc = [], 0
    for m in fs.models:
        bounds_base.append(acc)
        acc += m["bounds"].shape[0]
    t["bounds"] = np.concatenate([m["bounds"] for m in fs.models]).astype(f32) if fs.models else np.zeros((0, 6), f32)

    skinned_nodes = set()
    for node, b in fs.bones.items():
        mdl = fs.models[fs.model[node]]
        if len(b) > 0 and mdl["bind_pose"] is not None and len(b) == mdl["bind_pose"].shape[0]:
            skinned_nodes.add(node)
    skinned_mask = np.zeros(n, bool)
    if skinned_nodes:
        skinned_mask[list(skinned_nodes)] = True

    # rigid display entries: node index order x list order (scene.cpp:330-364); disabled nodes stay in the table with a flag
    ent_node, ent_list = [], []
    for mi, m in enumerate(fs.models):
        nodes = np.nonzero((fs.model == mi) & ~skinned_mask)[0]
        for li in range(m["bounds"].shape[0]):
            ok = nodes[m["mats"][li] < fs.n_mats[nodes]]  # mat_idx < materials.size() (scene.cpp:352)
            ent_node.append(ok)
            ent_list.append(np.full(len(ok), li, np.int64))
    ent_node = np.concatenate(ent_node) if ent_node else np.zeros(0, np.int64)
    ent_list = np.concatenate(ent_list) if ent_list else np.zeros(0, np.int64)
    order = np.lexsort((ent_list, ent_node))
    ent_node, ent_list = ent_node[order], ent_list[order]
    ent_model = fs.model[ent_node]
    mats_of = np.zeros(len(ent_node), np.int64)
    for mi, m in enumerate(fs.models):
        sel = ent_model == mi
        mats_of[sel] = m["mats"][ent_list[sel]]
    transparent = ((fs.transparent_mask[ent_node] >> mats_of.astype(np.uint32)) & 1).astype(bool)
    bidx = (np.array(bounds_base, np.int64)[ent_model] + ent_list) if len(ent_node) else np.zeros(0, np.int64)
    inst_of_node = None
    if fs.instance_start is not None and fs.instance_cameras is not None:
        inst_of_node = (np.searchsorted(fs.instance_start, np.arange(n), side="right") - 1).astype(np.uint16)
    for kind, sel in ((0, ~transparent), (1, transparent)):
        nodes_k = ent_node[sel]
        t["ent%d_node" % kind] = nodes_k.astype(np.uint32)
        t["ent%d_mtx_slot" % kind] = nodes_k.astype(np.uint32)
        t["ent%d_bounds_idx" % kind] = bidx[sel].astype(np.uint32)
        t["ent%d_mdl" % kind] = ent_model[sel].astype(np.uint16)
        t["ent%d_lst" % kind] = ent_list[sel].astype(np.uint16)
        t["ent%d_flags" % kind] = (fs.disabled[nodes_k] * ENTRY_DISABLED).astype(np.uint8)
        t["ent%d_view" % kind] = inst_of_node[nodes_k] if inst_of_node is not None else None
    if inst_of_node is not None:
        t["inst_entry_start"] = np.searchsorted(t["ent0_node"], fs.instance_start).astype(np.uint32)

    # skin records (scene.cpp:365-394), list by list in node order
    sk_mtx, sk_bind, sk_lists, bind_base, pool = [], [], [], {}, []
    acc = 0
    for mi, m in enumerate(fs.models):
        if m["bind_pose"] is not None:
            bind_base[mi] = acc
            acc += m["bind_pose"].shape[0]
            pool.append(m["bind_pose"])
    for node in sorted(skinned_nodes):
        if fs.disabled[node]:
            continue
        mi = int(fs.model[node])
        m, b = fs.models[mi], fs.bones[node]
        for li, table in enumerate(m["bone_tables"]):
            if m["mats"][li] >= fs.n_mats[node]:
                continue
            tbl = np.asarray(table, np.int64)
            slots = np.where(b[tbl] >= 0, b[tbl], node)
            sk_mtx.append(slots.astype(np.uint32))
            sk_bind.append((bind_base[mi] + tbl).astype(np.uint32))
            sk_lists.append((node, mi, li, len(tbl)))
    t["skin_mtx_slot"] = np.concatenate(sk_mtx) if sk_mtx else np.zeros(0, np.uint32)
    t["skin_bind_idx"] = np.concatenate(sk_bind) if sk_bind else np.zeros(0, np.uint32)
    t["skin_lists"] = sk_lists
    t["bind_pose"] = np.concatenate(pool).astype(f32) if pool else np.zeros((0, 12), f32)

    # bound tracks: for every play -> scene anim -> anims -> transform tracks
    tracks = []
    play_of_anim = np.full(len(fs.anim_node), -1, np.int64)
    for pi, (sa, _loop, _scale) in enumerate(fs.plays):
        play_of_anim[fs.scene_anims[sa][0]] = pi
    if len(fs.trk_anim):
        a = fs.trk_anim
        uq = fs.anim_use_quat[a].astype(bool)
        # AF_UseQuaternionForRotation picks the quaternion track and ignores the Euler one, and vice versa (scene.cpp:2168-2177)
        live = (play_of_anim[a] >= 0) & ~((fs.trk_kind == 1) & uq) & ~((fs.trk_kind == 3) & ~uq)
        rec = np.zeros(int(live.sum()), TRACK_DTYPE)
        rec["target_slot"] = fs.anim_node[a[live]]
        rec["key_off"] = fs.trk_key_off[live]
        rec["key_count"] = fs.trk_key_count[live]
        rec["play"] = play_of_anim[a[live]]
        rec["kind"] = fs.trk_kind[live]
        tracks = rec
    t["tracks"] = tracks if len(tracks) else np.zeros(0, TRACK_DTYPE)
    return t


def upload(ctx, fs, tables=None, fused_views=True):
    """push a FlatScene into a Context through the C ABI"""
    t = build_tables(fs) if tables is None else tables
    ctx.set_topology(t["parent_slot"])
    ctx.upload_trs(np.ascontiguousarray(fs.trs, f32))
    ctx.set_bounds(np.ascontiguousarray(t["bounds"]))
    for kind in (0, 1):
        ctx.set_display_entries(kind, t["ent%d_mtx_slot" % kind], t["ent%d_bounds_idx" % kind], t["ent%d_flags" % kind], t["ent%d_view" % kind])
    if len(t["skin_mtx_slot"]):
        ctx.set_skins(t["skin_mtx_slot"], t["skin_bind_idx"], t["bind_pose"])
    if len(t["tracks"]):
        ctx.set_anim_keys(fs.v3_t, np.ascontiguousarray(fs.v3_val), fs.v3_tension, fs.v3_bias, fs.q_t, np.ascontiguousarray(fs.q_val))
        ctx.set_tracks(t["tracks"], len(fs.plays))
    fr = scene_frusta(fs)
    if fs.instance_cameras is not None:
        ctx.set_instance_views(fr, t["inst_entry_start"])
    else:
        ctx.set_views(fr)
    return t


class PlayClocks:
    """Scene::PlayAnim + the clock half of UpdatePlayingAnims (scene.cpp:2353-2375, 2386-2413), exact int64 arithmetic."""

    def __init__(self, fs):
        self.t, self.t_start, self.t_end, self.t_scale, self.loop, self.alive = [], [], [], [], [], []
        for sa, loop, scale in fs.plays:
            _, ts, te = fs.scene_anims[sa]
            self.t.append(int(ts)); self.t_start.append(int(ts)); self.t_end.append(int(te))
            self.t_scale.append(int(np.int8(int(f32(scale) * f32(16.0))))); self.loop.append(loop); self.alive.append(True)
        self._remove = []

    def step(self, dt):
        """advance every clock; returns the int64 evaluation times (NOT_PLAYING for finished anims)"""
        for i in self._remove:  # ALM_Once anims are removed after their last evaluation
            self.alive[i] = False
        self._remove = []
        out = np.full(len(self.t), NOT_PLAYING, np.int64)
        for i in range(len(self.t)):
            if not self.alive[i]:
                continue
            t, sc = self.t[i] + ((dt * self.t_scale[i]) >> 4), self.t_scale[i]
            span = self.t_end[i] - self.t_start[i]
            if self.loop[i] == 2:
                if sc >= 0:
                    while t >= self.t_end[i]:
                        t -= span
                else:
                    while t <= self.t_start[i]:
                        t += span
            elif self.loop[i] == 0:
                if sc >= 0 and t >= self.t_end[i]:
                    t = self.t_end[i]; self._remove.append(i)
                elif sc < 0 and t <= self.t_start[i]:
                    t = self.t_start[i]; self._remove.append(i)
            self.t[i] = t
            out[i] = t
        return out
