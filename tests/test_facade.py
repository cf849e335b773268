"""The hg::Scene facade (harfang3d_b200/csrc/facade) driven through the SAME flat driver that drives the reference
(harfang3d_b200/csrc/facade/flat/hg_flat.cpp compiled twice: against /root/reference -> oracle/_ref/libhg_ref.so, against the facade
headers -> harfang3d_b200/libhg_facade_flat.so).

CPU tests: the host-side half of the facade — slot containers, refs, GetModelDisplayLists — needs no GPU and must match the
reference's behaviour (harfang/tests/engine/scene.cpp:115-154, foundation/vector_list.h).
GPU tests: whole frames (Update, display lists, CullModelDisplayLists, bone matrices, presets, edits) == the reference / golden fixtures.
"""
import ctypes as C
import os

import numpy as np
import pytest

from harfang3d_b200 import hostmirror, scenegen
from tests import oracle_lib
from tests.golden.gen_golden import golden_scenes
from tests.oracle_lib import P, bits
from tests.scene_io import RefScene

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def fac():
    L = oracle_lib.load_facade_driver()
    if L is None:
        pytest.skip("harfang3d_b200/libhg_facade_flat.so not built (make -C harfang3d_b200 facade)")
    return L


def six_objects():
    n = 6
    return scenegen.FlatScene(trs=np.tile(np.array([0, 0, 0, 0, 0, 0, 1, 1, 1], np.float32), (n, 1)), parent=np.full(n, -1, np.int32),
                              disabled=np.zeros(n, np.uint8), model=np.zeros(n, np.int32), n_mats=np.ones(n, np.uint8),
                              transparent_mask=np.zeros(n, np.uint32), models=scenegen.box_models(np.random.default_rng(0), 1),
                              camera=dict(scenegen.DEFAULT_CAMERA))


def test_facade_display_list_counts_like_reference_unit_test(fac):
    """harfang/tests/engine/scene.cpp:115-154 restated against the facade: 6 / 3 / 6 / 0 opaque lists (host-side, no GPU needed)"""
    sc = RefScene(fac, six_objects())
    counts = lambda: sc.display_lists()[0][0]
    assert counts() == 6
    for i in (0, 2, 4):
        fac.hgref_enable_node(sc.h, i, 0)
    assert counts() == 3
    for i in range(6):
        fac.hgref_enable_node(sc.h, i, 1)
    assert counts() == 6
    for i in range(6):
        fac.hgref_enable_node(sc.h, i, 0)
    assert counts() == 0
    sc.close()


@pytest.mark.ref
def test_facade_camera_component_host_side(fac, ref):
    """Camera component bookkeeping without a device: fov get / set, replaced components collected by GarbageCollect like in the reference,
    and the neutral (all-zero) view state for a missing current camera / a node without camera (scene.cpp:240-263 warn paths)"""
    fs = six_objects()
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    for L_, s in ((ref, a), (fac, b)):
        for x in s.view_state(16, 9):  # no current camera
            assert (bits(x) == 0).all()
        assert s.set_camera(2, 0.1, 100.0, fov=0.9) and s.set_camera(2, 0.1, 100.0, size=4.0, current=False)  # the first one is orphaned
        assert s.set_camera(4, 0.5, 50.0, fov=1.2, current=False)
        L_.hgref_set_camera_fov(s.h, 4, 0.33)
        assert L_.hgref_get_camera_fov(s.h, 4) == np.float32(0.33)
        L_.hgref_set_current_camera(s.h, 1)  # a node without a camera component
        for x in s.view_state(16, 9):
            assert (bits(x) == 0).all()
    assert ref.hgref_garbage_collect(a.h) == fac.hgref_garbage_collect(b.h) == 1
    assert ref.hgref_get_camera_fov(a.h, 4) == fac.hgref_get_camera_fov(b.h, 4) == np.float32(0.33)
    a.close(); b.close()


@pytest.mark.ref
def test_facade_slot_reuse_and_lists_match_reference(fac, ref):
    """destroy nodes + garbage collect + create again: transform slots are reused in the reference's order (LIFO free list), so the
    mtx_idx sequence of the display lists is identical (host-side)"""
    rng = np.random.default_rng(5)
    fs = scenegen.static_scene(400, 4, 6, n_models=3, seed=5, disabled_frac=0.1)
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    kill = rng.choice(fs.n, 120, replace=False)
    for L_, s in ((ref, a), (fac, b)):
        for nd in kill:
            L_.hgref_destroy_node(s.h, int(nd))
        assert L_.hgref_garbage_collect(s.h) > 0
    extra = scenegen.static_scene(90, 2, 89, n_models=3, seed=6, disabled_frac=0.0)
    for L_, s in ((ref, a), (fac, b)):
        L_.hgref_add_nodes(s.h, extra.n, P(np.ascontiguousarray(extra.trs)), None, P(extra.disabled), P(extra.model), P(extra.n_mats),
                           P(extra.transparent_mask), None)
    idx_a = [ref.hgref_node_transform_idx(a.h, i) for i in range(fs.n + extra.n)]
    idx_b = [fac.hgref_node_transform_idx(b.h, i) for i in range(fs.n + extra.n)]
    assert idx_a == idx_b
    (ca, la), (cb, lb) = a.display_lists(), b.display_lists()
    assert (ca == cb).all() and all((x == y).all() for x, y in zip(la[0], lb[0]))
    a.close(); b.close()


# ---- GPU: whole frames through the facade ---------------------------------------------------------------------------------------
def frame_checks(sc, gs, name, fs, frames, dt):
    for _ in range(frames):
        sc.update(dt)
    assert (bits(sc.worlds()[: fs.n]) == bits(gs[name + "_worlds"])).all()
    if frames > 1:
        assert (bits(sc.worlds(previous=True)[: fs.n]) == bits(gs[name + "_worlds_prev"])).all()
    counts, lists = sc.display_lists()
    assert (counts == gs[name + "_dl_counts"]).all() and (lists[0][0] == gs[name + "_dl_mtx"]).all()
    if name + "_trs" in gs:
        assert (bits(sc.trs(gs[name + "_trs_nodes"])) == bits(gs[name + "_trs"])).all()
    frusta = hostmirror.scene_frusta(fs)
    for v in range(len(frusta)):
        pos, mtx = sc.cull(0, frusta[v])
        assert (pos == gs["%s_vis%d_pos" % (name, v)]).all() and (mtx == gs["%s_vis%d_mtx" % (name, v)]).all()
    if name + "_bones" in gs:
        assert (bits(sc.bone_matrices()) == bits(gs[name + "_bones"])).all()
        assert (bits(sc.bone_matrices(previous=True)) == bits(gs[name + "_bones_prev"])).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c1", "pos", "c5s"])
def test_facade_frames_match_golden(fac, name):
    """Scene::Update + GetModelDisplayLists + CullModelDisplayLists through the facade == reference outputs (bit-exact)"""
    gs = np.load(os.path.join(G, "scenes.npz"))
    make, frames, dt = golden_scenes()[name]
    fs = make()
    sc = RefScene(fac, fs)
    frame_checks(sc, gs, name, fs, frames, dt)
    sc.close()


@pytest.mark.gpu
def test_facade_skinned_animated_within_tolerance(fac):
    """c3s has quaternion rotation tracks (device atan2): matrices within 1e-5 of the reference fixtures, lists identical"""
    gs = np.load(os.path.join(G, "scenes.npz"))
    make, frames, dt = golden_scenes()["c3s"]
    fs = make()
    sc = RefScene(fac, fs)
    for _ in range(frames):
        sc.update(dt)
    w, w_ref = sc.worlds()[: fs.n], gs["c3s_worlds"]
    scale = np.abs(w_ref).max()
    assert (np.abs(w.astype(np.float64) - w_ref) / np.maximum(np.abs(w_ref), 1e-6 * scale)).max() <= 1e-5
    b, b_ref = sc.bone_matrices(), gs["c3s_bones"]
    assert b.shape == b_ref.shape and (np.abs(b.astype(np.float64) - b_ref) / np.maximum(np.abs(b_ref), 1e-6 * np.abs(b_ref).max())).max() <= 1e-5
    counts, _ = sc.display_lists()
    assert (counts == gs["c3s_dl_counts"]).all()
    sc.close()


@pytest.mark.gpu
@pytest.mark.ref
def test_facade_presets_and_edits_match_reference(fac, ref):
    """SceneUpdateSystems-style split frame with SetNodeWorldMatrix presets, then TRS edits, re-parenting, enable/disable"""
    rng = np.random.default_rng(81)
    fs = scenegen.static_scene(3_000, 5, 6, n_models=2, seed=81)
    nodes = np.sort(rng.choice(fs.n, 700, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=10)
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    fr = hostmirror.scene_frusta(fs)[0]
    slots = rng.choice(fs.n, 150, replace=False)
    mats = np.zeros((150, 12), np.float32); mats[:, [0, 5, 10]] = 1; mats[:, [3, 7, 11]] = rng.uniform(-10, 10, (150, 3))

    def same():
        assert (bits(a.worlds()[: fs.n]) == bits(b.worlds()[: fs.n])).all()
        assert (bits(a.trs(np.arange(0, fs.n, 7))) == bits(b.trs(np.arange(0, fs.n, 7)))).all()
        (pa, ma), (pb, mb) = a.cull(0, fr), b.cull(0, fr)
        assert (pa == pb).all() and (ma == mb).all()

    for frame in range(3):
        for L_, s in ((ref, a), (fac, b)):
            L_.hgref_update_phase(s.h, 0, scenegen.FRAME_DT)
            if frame == 1:
                for sl, m in zip(slots, mats):
                    L_.hgref_set_node_world(s.h, int(sl), P(np.ascontiguousarray(m)))
            L_.hgref_update_phase(s.h, 1, scenegen.FRAME_DT)
        same()
    leaves = np.setdiff1d(np.arange(fs.n), np.unique(fs.parent[fs.parent >= 0]))
    moved = rng.choice(leaves, 200, replace=False)
    for L_, s in ((ref, a), (fac, b)):
        for i in range(0, fs.n, 3):
            t = fs.trs[i].copy(); t[3:6] += 0.1; t[0:3] -= 0.2
            L_.hgref_set_trs(s.h, i, P(np.ascontiguousarray(t)))
        for k, nd in enumerate(moved):
            L_.hgref_set_parent(s.h, int(nd), int(k % 30))
        for nd in range(0, fs.n, 11):
            L_.hgref_enable_node(s.h, nd, 0)
    for frame in range(2):
        a.update(scenegen.FRAME_DT); b.update(scenegen.FRAME_DT)
        same()
    a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.ref
@pytest.mark.parametrize("local", [False, True])
def test_facade_set_transform_matrices_match_reference(fac, ref, local):
    """Scene::SetTransformWorldMatrix / SetTransformLocalMatrix through the facade (queued, one batched device call): world rows
    bit-identical to the reference right after the calls, decomposed TRS within 1e-5, and the following frames agree."""
    rng = np.random.default_rng(91)
    fs = scenegen.static_scene(2_000, 4, 6, n_models=2, seed=91)
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    a.update(0); b.update(0)
    kids = rng.choice(np.nonzero(fs.parent >= 0)[0], 120, replace=False)
    order = np.concatenate([np.stack([kids, fs.parent[kids], kids], 1).reshape(-1), rng.choice(fs.n, 200, replace=False)])
    mats = np.zeros((len(order), 12), np.float32)
    trs = np.zeros((len(order), 9), np.float32)
    trs[:, 0:3] = rng.uniform(-20, 20, (len(order), 3)); trs[:, 3:6] = rng.uniform(-3, 3, (len(order), 3)); trs[:, 6:9] = rng.uniform(0.3, 2.5, (len(order), 3))
    trs[::5, 6] *= -1
    for i in range(len(order)):
        ref.hgref_transformation_mat4(P(trs[i]), P(mats[i]))
    for s in (a, b):
        for nd, m in zip(order, mats):
            s.set_transform_matrix(int(nd), m, local)
    assert (bits(a.worlds()[: fs.n]) == bits(b.worlds()[: fs.n])).all()
    touched = np.unique(order)
    ta, tb = a.trs(touched), b.trs(touched)
    assert (np.abs(ta.astype(np.float64) - tb) / np.maximum(np.abs(ta), 1e-6 * np.abs(ta).max())).max() <= 1e-5
    # an edit after the queued sets lands after them, like in the reference
    for s in (a, b):
        t = np.ascontiguousarray(fs.trs[int(order[0])], np.float32)
        s.L.hgref_set_trs(s.h, int(order[0]), P(t))
    for frame in range(2):
        a.update(scenegen.FRAME_DT); b.update(scenegen.FRAME_DT)
        wa, wb = a.worlds()[: fs.n].astype(np.float64), b.worlds()[: fs.n]
        assert (np.abs(wa - wb) / np.maximum(np.abs(wa), 1e-6 * np.abs(wa).max())).max() <= 1e-4
    a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.ref
def test_facade_scene_minmax_matches_reference(fac, ref):
    """Scene::GetMinMax / GetNodesMinMax through the facade == the reference, exactly (device reduction over the current worlds)"""
    fs = scenegen.static_scene(4_000, 4, 7, n_models=3, seed=95)
    fs.model[::9] = -1
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    a.update(0); b.update(0)
    for first, count in ((0, -1), (100, 900), (0, 1), (3990, 10)):
        (oka, ma), (okb, mb) = a.minmax(first, count), b.minmax(first, count)
        assert oka == okb and (not oka or (ma == mb).all()), (first, count, ma, mb)
    none = np.nonzero(fs.model < 0)[0]
    (oka, _), (okb, _) = a.minmax(int(none[0]), 1), b.minmax(int(none[0]), 1)
    assert not oka and not okb
    a.close(); b.close()


# ---- scene binary (format version 9) -----------------------------------------------------------------------------------------------------
def _binary_scene(kind):
    """scenes that exercise every section the loader keeps: hierarchy, disabled nodes, nodes without objects, transparent materials,
    multi-list models, skinned objects with bone refs, Hermite and quaternion tracks, scene anims"""
    if kind == "characters":
        return scenegen.config_c3(seed=33, n_chars=12, n_bones=64, n_keys=8)
    rng = np.random.default_rng(97)
    fs = scenegen.static_scene(2_500, 5, 6, n_models=3, seed=97)
    fs.model[rng.choice(fs.n, 200, replace=False)] = -1
    fs.transparent_mask[rng.choice(fs.n, 300, replace=False)] = 1
    scenegen.add_position_tracks(fs, np.sort(rng.choice(fs.n, 600, replace=False)).astype(np.int32), rng, n_keys=7)
    return fs


def _lists_equal(a, b):
    (ca, la), (cb, lb) = a.display_lists(), b.display_lists()
    assert (ca == cb).all(), (ca, cb)
    for k in (0, 1):
        for x, y in zip(la[k], lb[k]):
            assert (x == y).all()
    for kind in (2, 3):
        for x, y in zip(a.skinned_lists(kind), b.skinned_lists(kind)):
            assert x.shape == y.shape and (x == y).all()


@pytest.mark.ref
@pytest.mark.parametrize("kind", ["props", "characters"])
def test_scene_binary_loader_host_side(fac, ref, kind, tmp_path):
    """A file written by the reference's own Scene::Save_binary, read by (b) the reference's Load_binary and (c) the facade's bulk loader:
    same nodes, TRS, display lists (opaque / transparent / skinned incl. resolved bone indices). No GPU needed for this half."""
    fs = _binary_scene(kind)
    a = RefScene(ref, fs)
    path = str(tmp_path / "scene.scn")
    a.save_binary(path)
    assert os.path.getsize(path) > fs.n * 40
    b, c = RefScene(ref, fs, binary=path), RefScene(fac, fs, binary=path)
    every = np.arange(fs.n)
    ta, tb, tc = a.trs(every), b.trs(every), c.trs(every)
    assert (bits(ta) == bits(tb)).all() and (bits(tb) == bits(tc)).all()
    _lists_equal(a, b)
    _lists_equal(b, c)
    for s in (a, b, c):
        s.close()


@pytest.mark.ref
def test_scene_binary_loader_rejects_garbage(fac, tmp_path):
    """wrong magic, wrong version and truncated files are refused with a warning, never a crash"""
    ref = oracle_lib.load_ref()
    if ref is None:
        pytest.skip("needs the compiled reference to write a valid file")
    fs = _binary_scene("props")
    a = RefScene(ref, fs)
    path = str(tmp_path / "scene.scn")
    a.save_binary(path)
    a.close()
    blob = open(path, "rb").read()
    cases = {"magic": b"XXXX" + blob[4:], "version": blob[:5] + (8).to_bytes(4, "little") + blob[9:], "cut": blob[: len(blob) // 3], "empty": b""}
    for name, data in cases.items():
        p = str(tmp_path / (name + ".scn"))
        open(p, "wb").write(data)
        h = C.c_void_p(fac.hgref_create())
        assert fac.hgref_load_binary(h, os.fsencode(p)) == -1, name
        fac.hgref_destroy(h)
    h = C.c_void_p(fac.hgref_create())
    assert fac.hgref_load_binary(h, os.fsencode(str(tmp_path / "missing.scn"))) == -1
    fac.hgref_destroy(h)


@pytest.mark.gpu
@pytest.mark.ref
@pytest.mark.parametrize("kind", ["props", "characters"])
def test_scene_binary_loader_frames_match_reference(fac, ref, kind, tmp_path):
    """frames of the facade scene loaded in bulk from the binary == frames of the reference scene loaded by Load_binary from the same file"""
    fs = _binary_scene(kind)
    a = RefScene(ref, fs)
    path = str(tmp_path / "scene.scn")
    cams = []
    if kind == "props":  # camera components, the current camera and a "Camera.Fov" float track travel through the file too
        cams = [int(fs.anim_node[9]), int(fs.anim_node[5])]
        assert a.set_camera(cams[0], 0.5, 80.0, size=20.0, current=False) and a.set_camera(cams[1], 0.2, 250.0, fov=0.7)
        kt = np.arange(4, dtype=np.int64) * int(scenegen.KEY_DT * 0.8)
        ref.hgref_anim_add_float_track(a.h, 5, b"Camera.Fov", 4, P(kt), P(np.array([0.5, 1.4, 0.9, 2.0], np.float32)), P(np.array([0.2, -0.3, 0.1, 0], np.float32)),
                                       P(np.array([-0.1, 0.4, 0, 0.2], np.float32)))
    a.save_binary(path)
    a.close()
    b, c = RefScene(ref, fs, binary=path), RefScene(fac, fs, binary=path)
    fr = hostmirror.scene_frusta(fs)[0]
    # "Ready Matrices" at the end of the load: worlds exist before the first Update
    wb, wc = b.worlds()[: fs.n], c.worlds()[: fs.n]
    assert (bits(wb) == bits(wc)).all()
    for frame in range(4):
        b.update(scenegen.FRAME_DT); c.update(scenegen.FRAME_DT)
        wb, wc = b.worlds()[: fs.n].astype(np.float64), c.worlds()[: fs.n]
        assert (np.abs(wb - wc) / np.maximum(np.abs(wb), 1e-6 * np.abs(wb).max())).max() <= 1e-5
        if kind == "props":
            assert (bits(b.worlds()[: fs.n]) == bits(wc)).all()
        (pb, mb), (pc, mc) = b.cull(0, fr), c.cull(0, fr)
        assert (pb == pc).all() and (mb == mc).all()
        if cams:  # the current camera of the file, its fov driven by the loaded track
            fb, fc = ref.hgref_get_camera_fov(b.h, cams[1]), fac.hgref_get_camera_fov(c.h, cams[1])
            assert fb == fc and (frame == 0 or fb != np.float32(0.7)), (frame, fb, fc)
            for x, y in zip(b.view_state(16, 9), c.view_state(16, 9)):
                assert (bits(x) == bits(y)).all() and np.abs(x).max() > 0, frame
    if cams:  # the orthographic one
        ref.hgref_set_current_camera(b.h, cams[0]); fac.hgref_set_current_camera(c.h, cams[0])
        for x, y in zip(b.view_state(4, 3), c.view_state(4, 3)):
            assert (bits(x) == bits(y)).all() and np.abs(x).max() > 0
    if kind == "characters":
        bb, bc = b.bone_matrices().astype(np.float64), c.bone_matrices()
        assert bb.shape == bc.shape and (np.abs(bb - bc) / np.maximum(np.abs(bb), 1e-6 * np.abs(bb).max())).max() <= 1e-5
    b.close(); c.close()


@pytest.mark.gpu
@pytest.mark.ref
def test_facade_enable_tracks_match_reference(fac, ref):
    """"Enable" bool tracks through the facade: evaluated on the device, IsNodeEnabled / GetModelDisplayLists / CullModelDisplayLists follow
    the reference frame by frame; EnableNode / DisableNode from the host on tracked nodes after the anim ended behave like the reference."""
    from tests.test_gpu_parity import _enable_keys
    rng = np.random.default_rng(121)
    fs = scenegen.static_scene(2_000, 4, 6, n_models=2, seed=121)
    fs.model[rng.choice(fs.n, 150, replace=False)] = -1
    nodes = np.sort(rng.choice(fs.n, 500, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=6, play_loop=False)
    plays, fs.plays = fs.plays, []  # tracks first, PlayAnim after: the reference resolves an anim's tracks when it binds (BindNodeAnim)
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    fs.plays = plays
    tracked = np.arange(0, len(nodes), 3)
    kt, kv = _enable_keys(rng, len(tracked), 5, scenegen.KEY_DT)
    for L_, s in ((ref, a), (fac, b)):
        for k, an in enumerate(tracked):
            L_.hgref_anim_add_enable_track(s.h, int(an), 5, P(np.ascontiguousarray(kt[k])), P(np.ascontiguousarray(kv[k])))
        for sa, loop, scale in fs.plays:
            L_.hgref_play_anim(s.h, sa, loop, scale)
    fr = hostmirror.scene_frusta(fs)[0]
    watch = nodes[tracked]

    def same(frame):
        ea = [ref.hgref_is_node_enabled(a.h, int(n)) for n in watch]
        eb = [fac.hgref_is_node_enabled(b.h, int(n)) for n in watch]
        assert ea == eb, "frame %d" % frame
        (ca, la), (cb, lb) = a.display_lists(), b.display_lists()
        assert (ca == cb).all() and all((x == y).all() for x, y in zip(la[0], lb[0]))
        (pa, ma), (pb, mb) = a.cull(0, fr), b.cull(0, fr)
        assert (pa == pb).all() and (ma == mb).all()
        return sum(ea)

    counts = set()
    for frame in range(20):
        a.update(scenegen.FRAME_DT); b.update(scenegen.FRAME_DT)
        counts.add(same(frame))
    assert len(counts) > 3
    # the anim is over: host edits on tracked nodes stick
    for L_, s in ((ref, a), (fac, b)):
        for n in watch[:40]:
            L_.hgref_enable_node(s.h, int(n), 0)
        for n in watch[40:80]:
            L_.hgref_enable_node(s.h, int(n), 1)
    for frame in range(20, 23):
        a.update(scenegen.FRAME_DT); b.update(scenegen.FRAME_DT)
        same(frame)
    a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.ref
def test_facade_camera_view_state_matches_reference(fac, ref):
    """Camera component + Scene::ComputeCurrentCameraViewState (scene.cpp:240-263, render_pipeline.cpp:67-81): frustum planes, projection and
    view matrix bit-identical to the reference for a camera deep in the hierarchy (its world matrix comes from the device), perspective and
    orthographic, and the culls with those frusta keep the same lists; an invalid current camera yields the neutral view state on both."""
    fs = scenegen.static_scene(6_000, 5, 6, n_models=3, seed=123)
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    za, zb = a.view_state(16, 9), b.view_state(16, 9)  # no current camera: warn + {}
    for x, y in zip(za, zb):
        assert (bits(x) == bits(y)).all()
    depth = np.zeros(fs.n, np.int32)
    for i in range(fs.n):
        depth[i] = 0 if fs.parent[i] < 0 else depth[fs.parent[i]] + 1
    deep = np.nonzero(depth == depth.max())[0]
    cams = [(int(deep[0]), dict(fov=1.1)), (int(deep[-1]), dict(size=35.0)), (0, dict(fov=0.0005)), (int(deep[3]), dict(fov=3.5))]  # fov clamps: Deg(0.1), Deg(179.9)
    a.update(0); b.update(0)
    for frame in range(2):
        for node, kw in cams:
            for s in (a, b):
                assert s.set_camera(node, 0.1, 400.0, **kw)
            (fa, pa, va), (fb, pb, vb) = a.view_state(16, 9), b.view_state(16, 9)
            assert (bits(pa) == bits(pb)).all(), (node, kw)
            assert (bits(va) == bits(vb)).all(), (node, kw)
            assert (bits(fa) == bits(fb)).all(), (node, kw)
            for kind in (0, 1):
                (qa, ma), (qb, mb) = a.cull(kind, fa), b.cull(kind, fb)
                assert (qa == qb).all() and (ma == mb).all()
        t = np.ascontiguousarray(fs.trs[cams[0][0]] + np.float32(0.25), np.float32)  # move the first camera, next frame
        for s in (a, b):
            s.L.hgref_set_trs(s.h, cams[0][0], P(t))
            s.update(scenegen.FRAME_DT)
    assert a.L.hgref_garbage_collect(a.h) == b.L.hgref_garbage_collect(b.h)  # replaced camera components are collected alike
    a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.ref
def test_facade_camera_fov_track_matches_reference(fac, ref):
    """"Camera.Fov" float tracks (Hermite, BindNodeAnim scene.cpp:2091-2101 / EvaluateBoundAnim :2194-2197): evaluated on the host by the
    facade; the camera nodes also carry Position tracks sampled on the device. Fov and the whole view state of the current camera stay
    bit-identical to the reference frame by frame (before the first key, between keys, after the anim ended), the last track of that
    target wins and other float targets are ignored."""
    rng = np.random.default_rng(131)
    fs = scenegen.static_scene(3_000, 4, 6, n_models=2, seed=131)
    nodes = np.sort(rng.choice(np.nonzero(fs.parent >= 0)[0], 300, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=6, play_loop=False)
    plays, fs.plays = fs.plays, []
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    fs.plays = plays
    cam_anims = [3, 77, 200]
    kt = (np.arange(5, dtype=np.int64) * int(scenegen.KEY_DT * 1.3) + scenegen.KEY_DT // 3)
    tracks = []
    for _ in cam_anims:
        tracks.append([(tg, rng.uniform(0.3, 2.2, 5).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32))
                       for tg in (b"Camera.Fov", b"Light.DiffuseIntensity", b"Camera.Fov")])
    for L_, s in ((ref, a), (fac, b)):
        for an, trk in zip(cam_anims, tracks):
            assert s.set_camera(int(nodes[an]), 0.1, 300.0, fov=0.8, current=False)
            for tg, v, te, bi in trk:
                L_.hgref_anim_add_float_track(s.h, an, tg, 5, P(kt), P(v), P(te), P(bi))
        for sa, loop, scale in fs.plays:
            L_.hgref_play_anim(s.h, sa, loop, scale)
    seen = set()
    for frame in range(14):
        a.update(scenegen.FRAME_DT * 2); b.update(scenegen.FRAME_DT * 2)
        for an in cam_anims:
            fa, fb = ref.hgref_get_camera_fov(a.h, int(nodes[an])), fac.hgref_get_camera_fov(b.h, int(nodes[an]))
            assert np.float32(fa).view(np.uint32) == np.float32(fb).view(np.uint32), (frame, an, fa, fb)
            seen.add(fa)
            ref.hgref_set_current_camera(a.h, int(nodes[an])); fac.hgref_set_current_camera(b.h, int(nodes[an]))
            for x, y in zip(a.view_state(16, 9), b.view_state(16, 9)):
                assert (bits(x) == bits(y)).all(), (frame, an)
    assert len(seen) > 12  # the value really moves
    a.close(); b.close()


def test_reference_unit_tests_against_facade():
    """the reference's own scene unit tests (harfang/tests/engine/scene.cpp: ComponentGarbageCollection :27-33, WalkHierarchy :55-81,
    DisableObjectNodes :115-154), cut out of the reference source at build time and compiled UNCHANGED against the facade headers with the
    reference's test framework (oracle/facade_unit.cpp, `make -C oracle facade_unit`). Host only: no device call is reached."""
    import subprocess
    oracle_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle")
    exe = os.path.join(oracle_dir, "_ref", "facade_unit")
    if os.path.isdir("/root/reference/harfang"):
        subprocess.check_call(["make", "-C", oracle_dir, "facade_unit"], stdout=subprocess.DEVNULL)
    if not os.path.exists(exe):
        pytest.skip("needs the reference sources (or the prebuilt oracle/_ref/facade_unit)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    for name in ("ComponentGarbageCollection", "WalkHierarchy", "DisableObjectNodes"):
        assert ("engine.scene.%s" % name) in r.stdout and "[ OK ]" in r.stdout
    assert "All unit tests have passed" in r.stdout


@pytest.mark.gpu
def test_facade_light_and_material_tracks_match_reference(fac, ref):
    """the remaining single-value animation targets of BoundToNodeAnim (scene.h:74-88): "Light.Diffuse" / "Light.Specular" colour tracks,
    "Light.DiffuseIntensity" / "Light.SpecularIntensity" float tracks (EvaluateBoundAnim scene.cpp:2183-2192) and "Material.<slot>.<value>"
    vec4 tracks (:2199-2215), evaluated on the host by the facade while Position tracks of the same anims run on the device: every value
    bit-identical to the reference frame by frame, unknown slots / value names ignored like the reference ignores them."""
    rng = np.random.default_rng(141)
    fs = scenegen.static_scene(400, 3, 6, n_models=2, seed=141)
    nodes = np.sort(rng.choice(np.nonzero((fs.parent >= 0) & (fs.model >= 0))[0], 40, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=6, play_loop=True)
    plays, fs.plays = fs.plays, []
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    fs.plays = plays
    lit = [0, 7, 19]  # anim indices whose node gets a light + an animated material value
    kt = (np.arange(5, dtype=np.int64) * int(scenegen.KEY_DT * 0.9) + scenegen.KEY_DT // 4)
    col = [(tg, rng.uniform(0, 1, (5, 4)).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32))
           for tg in (b"Light.Diffuse", b"Light.Specular", b"Light.Diffuse")]
    flt = [(tg, rng.uniform(0.2, 4, 5).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32))
           for tg in (b"Light.DiffuseIntensity", b"Light.SpecularIntensity")]
    v4 = [(tg, rng.uniform(-2, 2, (5, 4)).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32), rng.uniform(-0.5, 0.5, 5).astype(np.float32))
          for tg in (b"Material.0.uColor", b"Material.3.uColor", b"Material.0.uMissing", b"Material.0.uColor")]
    start = np.array([0.1, 0.2, 0.3, 0.4], np.float32)
    for L_, s in ((ref, a), (fac, b)):
        for an in lit:
            nd = int(nodes[an])
            L_.hgref_add_point_light(s.h, nd, 2.5)
            L_.hgref_set_material_value(s.h, nd, 0, b"uColor", P(start))
            for tg, v, te, bi in col:
                L_.hgref_anim_add_color_track(s.h, an, tg, 5, P(kt), P(np.ascontiguousarray(v)), P(te), P(bi))
            for tg, v, te, bi in flt:
                L_.hgref_anim_add_float_track(s.h, an, tg, 5, P(kt), P(v), P(te), P(bi))
            for tg, v, te, bi in v4:
                L_.hgref_anim_add_vec4_track(s.h, an, tg, 5, P(kt), P(np.ascontiguousarray(v)), P(te), P(bi))
        for sa, loop, scale in fs.plays:
            L_.hgref_play_anim(s.h, sa, loop, scale)
    seen = set()
    for frame in range(16):
        a.update(scenegen.FRAME_DT * 2); b.update(scenegen.FRAME_DT * 2)
        assert (bits(a.worlds()[: fs.n]) == bits(b.worlds()[: fs.n])).all()
        for an in lit:
            la, lb, ma, mb = np.zeros(10, np.float32), np.zeros(10, np.float32), np.zeros(4, np.float32), np.zeros(4, np.float32)
            ref.hgref_get_light(a.h, int(nodes[an]), P(la)); fac.hgref_get_light(b.h, int(nodes[an]), P(lb))
            assert (bits(la) == bits(lb)).all(), (frame, an, la, lb)
            assert ref.hgref_get_material_value(a.h, int(nodes[an]), 0, b"uColor", P(ma)) == fac.hgref_get_material_value(b.h, int(nodes[an]), 0, b"uColor", P(mb)) == 1
            assert (bits(ma) == bits(mb)).all(), (frame, an, ma, mb)
            seen.add(float(la[0])); seen.add(float(ma[1]))
    assert len(seen) > 20  # the values really move
    a.close(); b.close()


@pytest.mark.gpu
def test_facade_scene_instances_match_reference(fac, ref, tmp_path):
    """Scene instances (scene.cpp:1304-1388): a node instantiates another scene file under itself (written here by the reference's own
    Save_binary). Checked against the reference through the shared flat binding: the instantiated nodes, their display lists and world
    matrices; NF_InstanceDisabled propagation when the host node is disabled / re-enabled (EnableNode_/DisableNode_ scene.cpp:472-503);
    the on-instantiate anim; and an instance_anim_track (scene.cpp:2217-2268) that switches the instantiated scene between two of its
    anims at a nested clock (looping, then reversed at half speed) — worlds bit-identical frame by frame."""
    rng = np.random.default_rng(151)
    # the instantiated scene: 60 nodes, two named scene anims driving different nodes
    sub = scenegen.static_scene(60, 3, 4, n_models=2, seed=152, disabled_frac=0.0)
    n_a = np.sort(rng.choice(np.arange(1, 60), 20, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(sub, n_a[:10], rng, n_keys=8, play_loop=True)
    scenegen.add_position_tracks(sub, n_a[10:], rng, n_keys=8, play_loop=True)
    sub.plays = []
    w = RefScene(ref, sub)
    ref.hgref_set_scene_anim_name(w.h, 0, b"idle")
    ref.hgref_set_scene_anim_name(w.h, 1, b"run")
    path = str(tmp_path / "sub.scn")
    w.save_binary(path)
    w.close()
    # the host scene: 12 nodes; nodes 3 and 7 instantiate the file (3 starts "idle" on instantiation), node 7's anim carries the step track
    fs = scenegen.static_scene(12, 2, 11, n_models=2, seed=153, disabled_frac=0.0)
    fs.models = sub.models
    a, b = RefScene(ref, fs), RefScene(fac, fs)
    kt = np.array([0, int(0.35e9), int(0.9e9)], np.int64)
    names = b"idle\0run\0idle\0"
    loop = np.array([2, 2, 0], np.uint8)
    scale = np.array([1.0, -0.5, 2.0], np.float32)
    for L_, s in ((ref, a), (fac, b)):
        assert L_.hgref_add_instance(s.h, 3, path.encode(), b"idle", 2) == 1
        assert L_.hgref_add_instance(s.h, 7, path.encode(), b"", 0) == 1
        assert L_.hgref_setup_instance(s.h, 3) == 60
        assert L_.hgref_setup_instance(s.h, 7) == 60
        assert L_.hgref_node_count(s.h) == 12 + 120
        an = L_.hgref_add_anim(s.h, 0)
        L_.hgref_anim_set_instance_track(s.h, an, 3, P(kt), names, P(loop), P(scale))
        sa = L_.hgref_add_scene_anim(s.h, 1, P(np.array([7], np.int32)), P(np.array([an], np.int32)), 0, int(1.4e9))
        L_.hgref_play_anim(s.h, sa, 0, 1.0)
    n_all = 132

    def same(frame):
        wa, wb = a.worlds()[:n_all], b.worlds()[:n_all]  # (the vectors are sized to the slot list's capacity, which grows differently)
        assert wa.shape == wb.shape == (n_all, 12) and (bits(wa) == bits(wb)).all(), "frame %s: worlds differ" % (frame,)
        (ca, la), (cb, lb) = a.display_lists(), b.display_lists()
        assert (ca == cb).all(), (frame, ca, cb)
        for k in (0, 1):
            for x, y in zip(la[k], lb[k]):
                assert (x == y).all()
        for nd in range(n_all):
            assert ref.hgref_is_node_enabled(a.h, nd) == fac.hgref_is_node_enabled(b.h, nd), (frame, nd)
        return int(ca[0])

    counts = []
    for frame in range(50):
        if frame == 10:
            ref.hgref_enable_node(a.h, 3, 0); fac.hgref_enable_node(b.h, 3, 0)   # the host of the first instance goes dark with all it instantiated
        if frame == 20:
            ref.hgref_enable_node(a.h, 3, 1); fac.hgref_enable_node(b.h, 3, 1)
        a.update(scenegen.FRAME_DT * 2); b.update(scenegen.FRAME_DT * 2)
        counts.append(same(frame))
    assert counts[15] < counts[5] and counts[25] == counts[5]  # the instantiated objects left and came back
    a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.ref
def test_facade_fused_lists_and_in_place_edits(fac, ref):
    """The fused path a HARFANG user calls — SetCullingFrustums, Scene::Update(dt), GetVisibleModelDisplayLists — returns the list
    GetModelDisplayLists + CullModelDisplayLists leave behind in the reference, frame after frame; with append slack, nodes spawned at the
    end of the live scene (CreateNode + CreateTransform + SetParent + CreateObject) and Enable / DisableNode reach the device in place:
    no depth sort, no table rebuild, no graph capture (launch count), same worlds and lists as the reference."""
    rng = np.random.default_rng(131)
    fs = scenegen.static_scene(6_000, 5, 6, n_models=2, seed=131)
    nodes = np.sort(rng.choice(fs.n, 1_500, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=10)
    fac.hgfacade_set_default_append_slack(1024)
    try:
        a, b = RefScene(ref, fs), RefScene(fac, fs)
    finally:
        fac.hgfacade_set_default_append_slack(0)
    fr = np.ascontiguousarray(hostmirror.scene_frusta(fs)[0])
    out3, n_all = np.zeros(3), [fs.n]

    def frame_and_compare():
        a.update(scenegen.FRAME_DT)
        nv = fac.hgfacade_time_fused_frames(b.h, 1, scenegen.FRAME_DT, P(fr), P(out3))
        got = np.zeros(max(nv, 1), np.uint32)
        assert fac.hgfacade_read_visible(b.h, P(got), len(got)) == nv
        _, mtx_ref = a.cull(0, fr)
        assert nv == len(mtx_ref) and (got[:nv] == mtx_ref).all()
        assert (bits(a.worlds()[: n_all[0]]) == bits(b.worlds()[: n_all[0]])).all()
        return nv

    for _ in range(5):  # (a caller that reads every frame gets the one-graph frame from the third frame on: both parities captured here)
        assert frame_and_compare() > 0
    # spawn 200 objects under random nodes, 40 of them under nodes of the same batch
    k = 200
    parent = rng.integers(0, fs.n, k).astype(np.int32)
    parent[160:] = fs.n + rng.integers(0, 160, 40)
    trs = scenegen.random_trs(rng, k, parent)
    trs[:, 6:9] = rng.uniform(0.8, 1.25, (k, 3))
    model = rng.integers(0, 2, k).astype(np.int32)
    for L_, s in ((ref, a), (fac, b)):
        L_.hgref_add_nodes(s.h, k, P(np.ascontiguousarray(trs)), P(parent), None, P(model), P(np.ones(k, np.uint8)), P(np.zeros(k, np.uint32)), None)
    n_all[0] = fs.n + k
    before = fac.hgfacade_rebuild_count(b.h)
    counts = []
    for _ in range(2):
        frame_and_compare()
        counts.append(fac.hgfacade_rebuild_count(b.h))
    for nd in list(rng.choice(fs.n + k, 30, replace=False)):  # flag flips on live tables
        for L_, s in ((ref, a), (fac, b)):
            L_.hgref_enable_node(s.h, int(nd), 0)
    frame_and_compare()
    counts.append(fac.hgfacade_rebuild_count(b.h))
    for nd in range(0, fs.n, 400):
        for L_, s in ((ref, a), (fac, b)):
            L_.hgref_enable_node(s.h, int(nd), 1)
    frame_and_compare()
    counts.append(fac.hgfacade_rebuild_count(b.h))
    assert counts == [before] * 4, "in-place edits must not re-sort, rebuild tables or capture graphs: %d -> %s" % (before, counts)
    a.close(); b.close()
