"""GPU parity tests: the CUDA path, called through the C ABI (libhgb200.so), against the oracle on the same seeded scenes.

Oracle = the UNMODIFIED reference compiled into oracle/_ref/libhg_ref.so when present (it travels to the GPU box prebuilt), else the
plain-C port, which is pinned bit-for-bit against it (tests/test_oracle_port_vs_ref.py, tests/golden).
Bars: visible lists / counts bit-exact; world and bone matrices within 1e-5 relative (north_star) — and in fact bit-identical for
every path that does not go through libm's double atan2 (quaternion rotation tracks), which is asserted separately.
"""
import ctypes as C

import numpy as np
import pytest

from harfang3d_b200 import hostmirror, scenegen
from tests import oracle_lib
from tests.oracle_lib import bits

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5  # north_star: world and bone matrices within 1e-5 relative


@pytest.fixture(scope="module")
def hgb():
    import harfang3d_b200
    return harfang3d_b200


@pytest.fixture(scope="module")
def oracles():
    return oracle_lib.load_ref(), oracle_lib.load_port()


def make_oracle(oracles, fs, tables):
    ref, port = oracles
    if ref is not None:
        from tests.scene_io import RefScene
        return RefScene(ref, fs)
    from tests.port_scene import PortScene
    return PortScene(port, fs, tables)


def rel_err(a, b):
    """max |a-b| / max(|b|, 1e-6*scale) with scale = max |b| of the whole array (SURVEY.md §8d parity gate)"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(float(np.abs(b).max()), 1e-30)
    return float((np.abs(a - b) / np.maximum(np.abs(b), 1e-6 * scale)).max())


def check_visible(ctx, t, orc, fr, kind=0, view=0):
    """GPU surviving entries -> (mtx_idx) sequence and position among ENABLED entries == reference's culled vector"""
    pos_ref, mtx_ref = orc.cull(kind, fr)
    idx = ctx.get_visible(kind, view)
    assert ctx.get_visible_count(kind, view) == len(idx)
    assert len(idx) == len(pos_ref), "visible count differs: gpu %d ref %d" % (len(idx), len(pos_ref))
    assert (np.diff(idx.astype(np.int64)) > 0).all(), "visible list must be strictly ascending (stable compaction)"
    flags = t["ent%d_flags" % kind]
    rank_enabled = np.cumsum(flags == 0) - 1
    assert (flags[idx] == 0).all()
    assert (rank_enabled[idx] == pos_ref).all()
    assert (t["ent%d_mtx_slot" % kind][idx] == mtx_ref).all()
    return len(idx)


def run_static(hgb, oracles, fs, fused, use_graph=True, expect_identity=None):
    ctx = hgb.Context(0, use_graph=use_graph)
    t = hostmirror.upload(ctx, fs)
    if expect_identity is not None:
        assert ctx.identity_order == expect_identity
    orc = make_oracle(oracles, fs, t)
    fr = hostmirror.scene_frusta(fs)[0]
    for frame in range(2):
        ctx.update(None, fused_cull=fused)
        if not fused:
            ctx.cull()
        orc.update(0)
        w = ctx.get_worlds()
        w_ref = orc.worlds()[: fs.n]
        assert rel_err(w, w_ref) <= REL_TOL
        assert (bits(w) == bits(w_ref)).all(), "static path is expected to be bit-identical, not just within tolerance"
        nvis = check_visible(ctx, t, orc, fr)
        assert 0 < nvis < len(t["ent0_flags"])
    # previous-frame worlds (a2: StorePreviousWorldMatrices is a buffer swap)
    assert (bits(ctx.get_worlds(previous=True)) == bits(orc.worlds(previous=True)[: fs.n])).all()
    orc.close(); ctx.close()
    return nvis


@pytest.mark.parametrize("fused", [False, True])
def test_c1_static_10k(hgb, oracles, fused):
    """BASELINE config 1: 10k nodes, 4 levels"""
    run_static(hgb, oracles, scenegen.config_c1(), fused, expect_identity=True)


def test_c1_plain_stream_launches(hgb, oracles):
    run_static(hgb, oracles, scenegen.config_c1(seed=11), True, use_graph=False)


def shuffled(fs, seed):
    """renumber nodes randomly: parents may come after children, slot order != depth order"""
    rng = np.random.default_rng(seed)
    perm = rng.permutation(fs.n)  # new index of old node i
    inv = np.argsort(perm)
    out = scenegen.FlatScene(trs=fs.trs[inv].copy(), parent=np.where(fs.parent[inv] >= 0, perm[np.maximum(fs.parent[inv], 0)], -1).astype(np.int32),
                             disabled=fs.disabled[inv].copy(), model=fs.model[inv].copy(), n_mats=fs.n_mats[inv].copy(),
                             transparent_mask=fs.transparent_mask[inv].copy(), models=fs.models, camera=fs.camera)
    return out


@pytest.mark.parametrize("fused", [False, True])
def test_shuffled_slot_order(hgb, oracles, fused):
    fs = shuffled(scenegen.static_scene(20_011, 6, 5, n_models=5, seed=21), 22)
    run_static(hgb, oracles, fs, fused, expect_identity=False)


def test_transparent_and_multi_list(hgb, oracles):
    rng = np.random.default_rng(31)
    fs = scenegen.static_scene(6_000, 5, 6, n_models=3, seed=31)
    # model 1 gets 3 lists using material slots 0,1,2; objects own 2 slots -> list 2 silently dropped (scene.cpp:352)
    lo = rng.uniform(-1, 0, (3, 3))
    fs.models[1] = dict(bounds=np.concatenate([lo, lo + rng.uniform(0.3, 2, (3, 3))], axis=1).astype(np.float32), mats=np.array([0, 1, 2], np.uint16),
                        bone_tables=None, bind_pose=None)
    fs.n_mats[:] = 2
    fs.transparent_mask[:] = rng.integers(0, 4, fs.n).astype(np.uint32)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = make_oracle(oracles, fs, t)
    fr = hostmirror.scene_frusta(fs)[0]
    ctx.update(None, fused_cull=True); orc.update(0)
    n0, n1 = check_visible(ctx, t, orc, fr, 0), check_visible(ctx, t, orc, fr, 1)
    assert n0 > 0 and n1 > 0
    # ComputeModelDisplayListSortKeys (scene_forward_pipeline.cpp:438-483) for the culled transparent list and the culled opaque list:
    # one 1-mm fixed-point key per survivor, bit-exact (the reference side runs over the list left by its last cull)
    if oracles[0] is not None:
        view = hostmirror.inverse_fast(np.asarray(fs.camera["world12"], np.float32).reshape(1, 12))[0]  # ViewState::view (render_pipeline.cpp:67-81)
        for kind, n_vis in ((1, n1), (0, n0)):
            orc.cull(kind, fr)
            keys_ref = orc.sort_keys(view)
            keys = ctx.sort_keys(kind, 0, view)
            assert len(keys) == n_vis == len(keys_ref)
            assert (keys == keys_ref).all(), "sort keys of list kind %d differ" % kind
        assert len(np.unique(keys)) > 10
    orc.close(); ctx.close()


def test_c2_static_1m(hgb, oracles):
    """BASELINE config 2: 1M static nodes, depth 8"""
    run_static(hgb, oracles, scenegen.config_c2(), True, expect_identity=True)


def test_edge_sizes(hgb, oracles):
    for n, depth, fan in [(1, 1, 1), (2, 2, 1), (33, 3, 4), (2048, 2, 2047), (2049, 3, 50), (4097, 12, 2)]:
        run_static_ok = scenegen.static_scene(n, depth, fan, n_models=2, seed=n, disabled_frac=0.0)
        run_static_ok.camera = dict(run_static_ok.camera, zfar=5000.0)
        ctx = hgb.Context(0)
        t = hostmirror.upload(ctx, run_static_ok)
        orc = make_oracle(oracles, run_static_ok, t)
        ctx.update(None, fused_cull=True); orc.update(0)
        assert (bits(ctx.get_worlds()) == bits(orc.worlds()[:n])).all()
        pos_ref, mtx_ref = orc.cull(0, hostmirror.scene_frusta(run_static_ok)[0])
        idx = ctx.get_visible(0, 0)
        assert (t["ent0_mtx_slot"][idx] == mtx_ref).all() and len(idx) == len(mtx_ref)
        orc.close(); ctx.close()


def test_empty_scene_and_no_entries(hgb):
    ctx = hgb.Context(0)
    ctx.set_topology(np.zeros(0, np.uint32))
    ctx.update(None, fused_cull=False)
    ctx.set_topology(np.full(5, 0xFFFFFFFF, np.uint32))
    ctx.upload_trs(np.tile(np.array([1, 2, 3, 0, 0, 0, 1, 1, 1], np.float32), (5, 1)))
    ctx.set_views(np.zeros((1, 24), np.float32))
    ctx.update(None, fused_cull=True)  # no display entries at all
    w = ctx.get_worlds()
    assert (w[:, [3, 7, 11]] == [1, 2, 3]).all()
    assert ctx.get_visible_count(0, 0) == 0
    ctx.close()


def test_errors_are_reported_not_thrown(hgb):
    ctx = hgb.Context(0)
    with pytest.raises(hgb.HgbError, match="cycle"):
        ctx.set_topology(np.array([1, 0], np.uint32))
    ctx.set_topology(np.array([0xFFFFFFFF, 0], np.uint32))
    with pytest.raises(hgb.HgbError, match="outside"):
        ctx.upload_trs(np.zeros((3, 9), np.float32))
    ctx.close()


# ---- animation, skins, presets, instance batches ----------------------------------------------------------------------------
def run_animated(hgb, oracles, fs, frames, dt=scenegen.FRAME_DT, with_bones=True, fused=True):
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = make_oracle(oracles, fs, t)
    clocks = hostmirror.PlayClocks(fs)
    fr = hostmirror.scene_frusta(fs)[0]
    has_quat = bool((fs.trk_kind == 3).any()) if len(fs.trk_kind) else False
    worst = 0.0
    for frame in range(frames):
        ctx.update(clocks.step(dt), fused_cull=fused)
        if not fused:
            ctx.cull()
        orc.update(dt)
        w, w_ref = ctx.get_worlds(), orc.worlds()[: fs.n]
        worst = max(worst, rel_err(w, w_ref))
        # every float of the path is bit-identical to the reference's: Hermite / linear evaluation is float-only arithmetic in the
        # reference's operation order, sinf / cosf are libm-exact, and the quaternion path's double atan2 is narrowed exactly
        # (hgb_atan2_exact.h) — so worlds, sampled TRS, visible lists and bone matrices are compared for EQUALITY, quaternion tracks included
        assert (bits(w) == bits(w_ref)).all(), "frame %d: world matrices differ from the oracle (quaternion tracks: %s)" % (frame, has_quat)
        animated = np.unique(fs.anim_node)
        trs_ref = orc.trs(animated[:: max(1, len(animated) // 500)])
        trs_gpu = ctx.get_trs()[animated[:: max(1, len(animated) // 500)]]
        assert (bits(trs_gpu) == bits(np.ascontiguousarray(trs_ref, np.float32))).all(), "frame %d: sampled TRS differ" % frame
        if len(t["ent0_flags"]):
            pos_ref, mtx_ref = orc.cull(0, fr)
            idx = ctx.get_visible(0, 0)
            assert len(idx) == len(mtx_ref) and (t["ent0_mtx_slot"][idx] == mtx_ref).all(), "frame %d: visible list differs" % frame
        if with_bones and len(t["skin_mtx_slot"]):
            ctx.bone_matrices(with_previous=frame > 0)
            b, b_ref = ctx.get_bone_matrices(), orc.bone_matrices()
            assert b.shape == b_ref.shape
            assert (bits(b) == bits(np.ascontiguousarray(b_ref, np.float32))).all(), "frame %d: bone matrices differ" % frame
            if frame > 0:
                assert (bits(ctx.get_bone_matrices(previous=True)) == bits(np.ascontiguousarray(orc.bone_matrices(previous=True), np.float32))).all()
    orc.close(); ctx.close()
    return worst


def test_c3_small_skinned_animated(hgb, oracles):
    """BASELINE config 3 at 48 characters x 64 bones: Hermite pos/rot/scale tracks, quaternion rotation tracks on odd characters,
    ALM_Loop clocks over 40 frames (wraps the 1.55 s loop once at dt = 50 ms)"""
    fs = scenegen.config_c3(n_chars=48, n_bones=64, n_keys=32)
    run_animated(hgb, oracles, fs, frames=40, dt=50_000_000)


def test_position_tracks_bit_exact(hgb, oracles):
    rng = np.random.default_rng(41)
    fs = scenegen.static_scene(30_000, 6, 6, n_models=4, seed=41)
    nodes = np.sort(rng.choice(fs.n, 9_000, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=16)
    run_animated(hgb, oracles, fs, frames=6, with_bones=False)


def test_play_once_ends_and_negative_scale(hgb, oracles):
    rng = np.random.default_rng(43)
    fs = scenegen.static_scene(3_000, 4, 8, n_models=2, seed=43)
    scenegen.add_position_tracks(fs, np.arange(0, 1000, dtype=np.int32), rng, n_keys=4, play_loop=False)   # ALM_Once, ends after 150 ms
    scenegen.add_position_tracks(fs, np.arange(1500, 2500, dtype=np.int32), rng, n_keys=8, play_loop=True)
    fs.plays[1] = (1, 2, -0.5)  # looped, played backwards at half speed
    run_animated(hgb, oracles, fs, frames=16, dt=33_333_333, with_bones=False)


def test_preset_world_matrices(hgb, oracles):
    """SetNodeWorldMatrix between ReadyWorldMatrices and ComputeWorldMatrices wins and is the parent of its children
    (scene.cpp:768-786; how physics/scripts inject transforms, scene_systems.cpp:153-179)"""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    from tests.oracle_lib import P
    fs = scenegen.static_scene(5_000, 5, 6, n_models=2, seed=51)
    rng = np.random.default_rng(52)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    slots = np.sort(rng.choice(fs.n, 300, replace=False)).astype(np.uint32)
    mats = np.zeros((300, 12), np.float32)
    mats[:, [0, 5, 10]] = 1
    mats[:, [3, 7, 11]] = rng.uniform(-20, 20, (300, 3))
    for frame in range(3):
        ref.hgref_update_phase(orc.h, 0, 0)
        if frame != 1:  # frame 1 has no presets: the mask must have been cleared
            for s, m in zip(slots, mats):
                ref.hgref_set_node_world(orc.h, int(s), P(np.ascontiguousarray(m)))
            ctx.set_preset_world(slots, mats)
        ref.hgref_update_phase(orc.h, 1, 0)
        ctx.update(None, fused_cull=True)
        assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all(), "frame %d" % frame
        check_visible(ctx, t, orc, hostmirror.scene_frusta(fs)[0])
    orc.close(); ctx.close()


def _random_affine(rng, n, negative_every=7):
    """world matrices = T * R(YXZ euler) * S with some negative and some tiny scales (Decompose's sign fold, matrix4.cpp:226-229)"""
    from tests import oracle_lib
    port = oracle_lib.load_port()
    trs = np.zeros((n, 9), np.float32)
    trs[:, 0:3] = rng.uniform(-30, 30, (n, 3))
    trs[:, 3:6] = rng.uniform(-np.pi, np.pi, (n, 3))
    trs[:, 6:9] = rng.uniform(0.25, 3.0, (n, 3))
    trs[::negative_every, 6] *= -1
    out = np.zeros((n, 12), np.float32)
    for i in range(n):
        port.hgp_transformation_mat4(oracle_lib.P(trs[i]), oracle_lib.P(out[i]))
    return out


@pytest.mark.parametrize("local", [False, True])
def test_set_transform_matrices_batched(hgb, oracles, local):
    """Scene::SetTransformWorldMatrix / SetTransformLocalMatrix (component.cpp:265-281, :236-248) as one batched device call: world rows
    bit-identical, the decomposed TRS within 1e-5 (Euler angles go through libm's double atan2), sequential semantics when a batch sets a
    node and its parent, and the next Update propagates from the refreshed TRS."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    fs = scenegen.static_scene(6_000, 5, 6, n_models=2, seed=61)
    rng = np.random.default_rng(62)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    ctx.update(None, fused_cull=True)
    orc.update(0)
    # 500 random nodes + 200 (child, parent, child-again) chains: read-after-write and write-after-write inside one batch
    picks = rng.choice(fs.n, 500, replace=False)
    kids = rng.choice(np.nonzero(fs.parent >= 0)[0], 200, replace=False)
    chain = np.stack([kids, fs.parent[kids], kids], 1).reshape(-1)
    slots = np.concatenate([picks, chain, picks[:50]]).astype(np.uint32)
    mats = _random_affine(rng, len(slots))
    for s, m in zip(slots, mats):
        orc.set_transform_matrix(int(s), m, local)
    ctx.set_matrices(slots, mats, local)
    w, w_ref = ctx.get_worlds(), orc.worlds()[: fs.n]
    assert (bits(w) == bits(w_ref)).all(), "world rows after the batched set differ"
    touched = np.unique(slots)
    trs, trs_ref = ctx.get_trs()[touched], orc.trs(touched)
    assert rel_err(trs, trs_ref) <= REL_TOL
    print("set_matrices(local=%s): decomposed TRS bit-identical on %d of %d values" % (local, int((bits(trs) == bits(trs_ref)).sum()), trs.size))
    untouched = np.setdiff1d(np.arange(fs.n), touched)
    assert (bits(ctx.get_trs()[untouched]) == bits(np.ascontiguousarray(fs.trs, np.float32)[untouched])).all(), "rows of other transforms were disturbed"
    # the next frame recomputes every world from the TRS (reference: Update -> ComputeWorldMatrices)
    ctx.update(None, fused_cull=True)
    orc.update(0)
    assert rel_err(ctx.get_worlds(), orc.worlds()[: fs.n]) <= 1e-4  # TRS within 1e-5, composed through up to 5 levels
    orc.close(); ctx.close()


def _same_box(a, b):
    """min/max are exact; only the sign of a zero may differ (Union's `a < b ? a : b` keeps whichever came first)"""
    return bool((np.asarray(a, np.float32) == np.asarray(b, np.float32)).all())


def test_scene_minmax_reduction(hgb, oracles):
    """Scene::GetMinMax (scene.cpp:1982) = union over object nodes of world * Union(model bounds): device reduction == reference, exactly,
    for a static frame, after track sampling, and after SetTransformWorldMatrix edits; GetNodesMinMax over a node range as one segment."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    fs = scenegen.static_scene(20_000, 5, 7, n_models=5, seed=71)
    rng = np.random.default_rng(72)
    fs.model[rng.choice(fs.n, 3000, replace=False)] = -1          # nodes without an Object do not take part
    fs.models[2]["bounds"] = np.concatenate([fs.models[2]["bounds"], fs.models[2]["bounds"] * np.float32(1.7)])  # a model with two lists
    fs.models[2]["mats"] = np.array([0, 0], np.uint16)
    scenegen.add_position_tracks(fs, np.sort(rng.choice(fs.n, 4000, replace=False)).astype(np.int32), rng, n_keys=6)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    slots, ob_idx, ob, seg, n_seg = hostmirror.object_table(fs)
    ctx.set_objects(slots, ob_idx, ob, seg, n_seg)
    clocks = hostmirror.PlayClocks(fs)
    for frame in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
        orc.update(scenegen.FRAME_DT)
        mm, cnt = ctx.minmax()
        ok, mm_ref = orc.minmax()
        assert ok and cnt[0] == len(slots)
        assert _same_box(mm[0], mm_ref), (mm, mm_ref)
    # an edit that pushes one node far out must show up at once (world rows are written by the batched set)
    far = np.array([[1, 0, 0, 900.0, 0, 1, 0, -700.0, 0, 0, 1, 1234.5]], np.float32)
    node = int(slots[17])
    ctx.set_matrices(np.array([node], np.uint32), far)
    orc.set_transform_matrix(node, far[0])
    mm, _ = ctx.minmax()
    ok, mm_ref = orc.minmax()
    assert ok and _same_box(mm[0], mm_ref) and mm[0][3] > 890
    # GetNodesMinMax over a node range = the same reduction over a subset
    lo, hi = 5000, 9000
    sel = (slots >= lo) & (slots < hi)
    ctx.set_objects(slots[sel], ob_idx[sel], ob, None, 1)
    mm, cnt = ctx.minmax()
    ok, mm_ref = orc.minmax(lo, hi - lo)
    assert ok and cnt[0] == int(sel.sum()) and _same_box(mm[0], mm_ref)
    # no object at all: GetNodesMinMax returns false
    ctx.set_objects(slots[:0], ob_idx[:0], ob, None, 1)
    _, cnt = ctx.minmax()
    assert cnt[0] == 0
    orc.close(); ctx.close()


def test_instance_minmax_segments(hgb, oracles):
    """one bounding box per scene instance of an RL-style batch (segmented reduction; instances interleave in position order)"""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    fs = scenegen.config_c5(seed=73, n_instances=48, nodes_per_instance=300, depth=4)
    fs.model[::13] = -1
    ctx = hgb.Context(0)
    hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    ctx.update(None, fused_cull=True)
    orc.update(0)
    slots, ob_idx, ob, seg, n_seg = hostmirror.object_table(fs)
    assert n_seg == 48
    ctx.set_objects(slots, ob_idx, ob, seg, n_seg)
    mm, cnt = ctx.minmax()
    for i in range(n_seg):
        lo, hi = int(fs.instance_start[i]), int(fs.instance_start[i + 1])
        ok, mm_ref = orc.minmax(lo, hi - lo)
        assert ok and cnt[i] == int(((slots >= lo) & (slots < hi)).sum())
        assert _same_box(mm[i], mm_ref), "instance %d" % i
    orc.close(); ctx.close()


def test_c5_small_instance_batch(hgb, oracles):
    """BASELINE config 5 at 24 instances x 2048 nodes: one camera per instance; per-instance visible lists = the reference's
    CullModelDisplayLists run per instance scene"""
    fs = scenegen.config_c5(n_instances=24)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    ctx.update(None, fused_cull=True)
    idx = ctx.get_visible(0, 0)
    off = ctx.get_instance_offsets()
    assert off[0] == 0 and off[-1] == len(idx) and (np.diff(off.astype(np.int64)) >= 0).all()
    frusta = hostmirror.scene_frusta(fs)
    orc = make_oracle(oracles, fs, t)  # one big reference scene holding all instances: worlds are per-node anyway
    orc.update(0)
    assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all()
    ent_node = t["ent0_node"]
    for i in range(len(frusta)):
        pos_ref, mtx_ref = orc.cull(0, frusta[i])
        lo, hi = fs.instance_start[i], fs.instance_start[i + 1]
        mine = mtx_ref[(mtx_ref >= lo) & (mtx_ref < hi)]
        got = ent_node[idx[off[i]:off[i + 1]]]
        assert len(got) == len(mine) and (got == mine).all(), "instance %d" % i
    # stand-alone cull gives the same lists
    ctx.cull()
    assert (ctx.get_visible(0, 0) == idx).all() and (ctx.get_instance_offsets() == off).all()
    orc.close(); ctx.close()


def test_multi_view_cull(hgb, oracles):
    """6 frusta in one pass (view + 4 PSSM splits + spot in the reference: forward_pipeline.cpp:331-332,380-381)"""
    fs = scenegen.static_scene(40_000, 6, 7, n_models=6, seed=61)
    rng = np.random.default_rng(62)
    cams = scenegen.random_camera_worlds(rng, 6, extent=10.0)
    c = fs.camera
    frusta = hostmirror.make_perspective_frustum(c["znear"], c["zfar"], c["fov"], c["ar_x"], c["ar_y"], cams)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    ctx.set_views(frusta)
    orc = make_oracle(oracles, fs, t)
    orc.update(0)
    for fused in (True, False):
        ctx.update(None, fused_cull=fused)
        if not fused:
            ctx.cull()
        for v in range(6):
            check_visible(ctx, t, orc, frusta[v], 0, v)
    orc.close(); ctx.close()


def test_profile_counters(hgb):
    fs = scenegen.config_c1()
    ctx = hgb.Context(0, top_walk=-1)
    hostmirror.upload(ctx, fs)
    ctx.set_profiling(True)
    ctx.update(None, fused_cull=True)
    prof = ctx.get_profile()
    assert prof["k_update_level"][1] == 4 and prof["k_update_level"][0] > 0 and prof["k_compact"][1] == 1  # one launch per level
    w4 = ctx.get_worlds()
    ctx.close()
    ctx = hgb.Context(0)  # default: the four small levels of the 10k-node scene as ONE launch (k_update_top, ancestor walk)
    hostmirror.upload(ctx, fs)
    ctx.set_profiling(True)
    ctx.update(None, fused_cull=True)
    assert ctx.get_profile()["k_update_level"][1] == 1
    assert (bits(ctx.get_worlds()) == bits(w4)).all()
    ctx.close()
    ctx = hgb.Context(0, top_walk=25_000)  # budget reached after three levels (1 + 21*2 + 441*3): top launch + one level launch
    hostmirror.upload(ctx, fs)
    ctx.set_profiling(True)
    ctx.update(None, fused_cull=True)
    assert ctx.get_profile()["k_update_level"][1] == 2
    assert (bits(ctx.get_worlds()) == bits(w4)).all()
    ctx.close()


def test_edits_while_animated(hgb, oracles):
    """TRS uploads, re-parenting and track re-binding on a scene whose Position triples are track-driven (the sampled values live in
    a side array on the device): every observable (worlds, TRS, visible list) keeps matching the reference's Scene after each edit."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    from tests.oracle_lib import P
    rng = np.random.default_rng(71)
    fs = scenegen.static_scene(4_000, 5, 6, n_models=3, seed=71)
    nodes = np.sort(rng.choice(fs.n, 1_200, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=12)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    clocks = hostmirror.PlayClocks(fs)
    fr = hostmirror.scene_frusta(fs)[0]

    def frame():
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
        orc.update(scenegen.FRAME_DT)
        assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all()
        assert (bits(ctx.get_trs()) == bits(orc.trs(np.arange(fs.n)))).all()
        check_visible(ctx, t, orc, fr)

    frame(); frame()
    # 1. the host rewrites every TRS (SetTransformTRS on animated nodes too: the track overwrites pos again next frame)
    new_trs = fs.trs.copy()
    new_trs[:, 3:6] += rng.uniform(-0.3, 0.3, (fs.n, 3)).astype(np.float32)
    new_trs[:, 0:3] += 0.25
    ctx.upload_trs(np.ascontiguousarray(new_trs))
    for i in range(fs.n):
        ref.hgref_set_trs(orc.h, i, P(np.ascontiguousarray(new_trs[i])))
    assert (bits(ctx.get_trs()) == bits(orc.trs(np.arange(fs.n)))).all()
    frame()
    # 2. re-parent a few hundred leaves under other nodes of a smaller index (no cycles): the depth-sorted layout is rebuilt
    parent = fs.parent.copy()
    leaves = np.setdiff1d(np.arange(fs.n), np.unique(parent[parent >= 0]))
    moved = rng.choice(leaves, 300, replace=False)
    for nd in moved:
        parent[nd] = int(rng.integers(0, 40))
        ref.hgref_set_parent(orc.h, int(nd), int(parent[nd]))
    fs.parent = parent
    ctx.set_topology(np.where(parent >= 0, parent, 0xFFFFFFFF).astype(np.uint32))
    frame(); frame()
    orc.close(); ctx.close()


def _enable_keys(rng, n_tracks, n_keys, key_dt):
    """bool step tracks: key times on a jittered grid (some before t=0, so 'before the first key' is exercised), random values"""
    t = (np.arange(n_keys, dtype=np.int64)[None, :] * key_dt + rng.integers(-key_dt // 3, key_dt // 3, (n_tracks, n_keys))).astype(np.int64)
    t[: n_tracks // 4, 0] += 2 * key_dt  # a quarter of the tracks start late
    t.sort(axis=1)
    v = rng.integers(0, 2, (n_tracks, n_keys)).astype(np.uint8)
    return t, v


def test_enable_tracks_on_device(hgb, oracles):
    """"Enable" bool tracks (EvaluateBoundAnim scene.cpp:2158-2162, EvaluateStep animation.h:130-139): nodes switch on and off from frame to
    frame; visible lists and the per-node state equal the reference's, the state persists after the anim ended (ALM_Once) and survives a
    table rebuild."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from harfang3d_b200.capi import TRACK_DTYPE
    from tests.oracle_lib import P
    from tests.scene_io import RefScene
    rng = np.random.default_rng(111)
    fs = scenegen.static_scene(4_000, 4, 7, n_models=3, seed=111)
    fs.model[rng.choice(fs.n, 400, replace=False)] = -1  # some tracked nodes own no display entry
    nodes = np.sort(rng.choice(fs.n, 900, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=6, play_loop=False)  # one anim per node, one scene anim, ALM_Once: ends after 5 key intervals
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    plays, fs.plays = fs.plays, []   # the reference resolves an anim's tracks when PlayAnim binds it (BindNodeAnim): tracks first, play after
    orc = RefScene(ref, fs)
    fs.plays = plays
    tracked = np.arange(0, len(nodes), 2)  # every second anim gets an Enable track
    kt, kv = _enable_keys(rng, len(tracked), 5, scenegen.KEY_DT)
    tr = np.zeros(len(tracked), TRACK_DTYPE)
    tr["target_slot"] = nodes[tracked]
    tr["key_off"] = np.arange(len(tracked)) * 5
    tr["key_count"] = 5
    tr["play"] = 0
    tr["kind"] = 4
    ctx.set_enable_tracks(tr, np.ascontiguousarray(kt.reshape(-1)), np.ascontiguousarray(kv.reshape(-1)))
    for k, a in enumerate(tracked):
        ref.hgref_anim_add_enable_track(orc.h, int(a), 5, P(np.ascontiguousarray(kt[k])), P(np.ascontiguousarray(kv[k])))
    for sa, loop, scale in fs.plays:
        ref.hgref_play_anim(orc.h, sa, loop, scale)
    fr = hostmirror.scene_frusta(fs)[0]
    clocks = hostmirror.PlayClocks(fs)
    seen = set()
    for frame in range(24):  # the anim (250 ms) ends around frame 15: the last frames check that the state sticks
        if frame == 9:        # a table rebuild in the middle (entry flags re-sent by the host): the tracked state must survive it
            ctx.set_entry_flags(0, 0, t["ent0_flags"][:1].copy())
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
        orc.update(scenegen.FRAME_DT)
        _, mtx_ref = orc.cull(0, fr)
        idx = ctx.get_visible(0, 0)
        assert len(idx) == len(mtx_ref) and (t["ent0_mtx_slot"][idx] == mtx_ref).all(), "frame %d" % frame
        slots, dis = ctx.get_enable_state()
        assert (slots == np.sort(nodes[tracked])).all()
        dis_ref = np.array([0 if ref.hgref_is_node_enabled(orc.h, int(s)) else 1 for s in slots], np.uint8)
        assert (dis == dis_ref).all(), "frame %d" % frame
        seen.add(int(dis.sum()))
        assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all()
    assert len(seen) > 3, "the tracks are expected to switch nodes on and off over the frames"
    orc.close(); ctx.close()


# ---- round 2: chain / edit / fetch regressions --------------------------------------------------------------------------------
def test_preset_root_with_animated_children_chained(hgb, oracles):
    """A single-root scene whose root is pre-set with SetNodeWorldMatrix while its descendants are track-driven: with the frame's
    kernels chained by programmatic dependent launch every thread of the first level takes the preset branch, which must still pass the
    dependency wait (otherwise later levels could read last frame's samples). Several frames, graph and plain launches."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    from tests.oracle_lib import P
    rng = np.random.default_rng(91)
    fs = scenegen.static_scene(6_000, 5, 8, n_models=2, seed=91)
    assert int((fs.parent < 0).sum()) == 1
    nodes = np.sort(rng.choice(np.arange(1, fs.n), 2_500, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=10)
    for top_walk in (0, -1):  # the one-launch top of the hierarchy and one launch per level
        ctx = hgb.Context(0, top_walk=top_walk)
        t = hostmirror.upload(ctx, fs)
        orc = RefScene(ref, fs)
        clocks = hostmirror.PlayClocks(fs)
        root = np.array([int(np.nonzero(fs.parent < 0)[0][0])], np.uint32)
        for frame in range(6):
            m = np.zeros((1, 12), np.float32)
            m[0, [0, 5, 10]] = 1
            m[0, [3, 7, 11]] = [frame * 0.5, 1.0, -2.0]
            ref.hgref_update_phase(orc.h, 0, int(scenegen.FRAME_DT))
            ref.hgref_set_node_world(orc.h, int(root[0]), P(np.ascontiguousarray(m[0])))
            ref.hgref_update_phase(orc.h, 1, int(scenegen.FRAME_DT))
            ctx.set_preset_world(root, m)
            ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
            assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all(), "top_walk %d frame %d" % (top_walk, frame)
            check_visible(ctx, t, orc, hostmirror.scene_frusta(fs)[0])
        orc.close(); ctx.close()


def test_previous_worlds_survive_hierarchy_edits(hgb, oracles):
    """transform_worlds / previous_transform_worlds are indexed by slot in the reference and survive SetParent / CreateNode
    (scene.cpp:280-304): after a re-parenting (which re-sorts the device layout) the previous-frame matrices — and the previous-frame
    bone twins that read them — still equal the reference's."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    rng = np.random.default_rng(92)
    fs = scenegen.static_scene(3_000, 5, 6, n_models=2, seed=92)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    fr = hostmirror.scene_frusta(fs)[0]
    for frame in range(2):
        ctx.update(None, fused_cull=True)
        orc.update(0)
    # move a batch of leaves under other parents (no cycles: leaves only), then one more frame
    leaves = np.setdiff1d(np.arange(fs.n), fs.parent[fs.parent >= 0])
    moved = rng.choice(leaves, 200, replace=False)
    new_parent = rng.choice(np.setdiff1d(np.arange(fs.n), leaves), 200)
    parent = fs.parent.copy()
    for n_, p_ in zip(moved, new_parent):
        parent[n_] = p_
        ref.hgref_set_parent(orc.h, int(n_), int(p_))
    fs.parent = parent
    t2 = hostmirror.build_tables(fs)
    trs_before = ctx.get_trs()
    ctx.set_topology(t2["parent_slot"])
    assert (bits(ctx.get_trs()) == bits(trs_before)).all()
    assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all(), "current worlds must survive the re-sort"
    assert (bits(ctx.get_worlds(previous=True)) == bits(orc.worlds(previous=True)[: fs.n])).all(), "previous worlds must survive the re-sort"
    ctx.update(None, fused_cull=True)
    orc.update(0)
    assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all()
    assert (bits(ctx.get_worlds(previous=True)) == bits(orc.worlds(previous=True)[: fs.n])).all(), "previous worlds after the edit"
    check_visible(ctx, t, orc, fr)
    orc.close(); ctx.close()


def test_async_tail_and_one_round_trip_fetch(hgb):
    """the asynchronous frame tail (compaction on a second stream, visibility bytes double-buffered) returns what the one-graph frame
    returns, frame after frame with edits in between; hgb_get_visible_ptr == hgb_get_visible, also when the list outgrows the size guess"""
    rng = np.random.default_rng(93)
    fs = scenegen.static_scene(60_000, 6, 7, n_models=3, seed=93)
    nodes = np.sort(rng.choice(fs.n, 20_000, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=12)
    a, b = hgb.Context(0, async_post=True), hgb.Context(0, async_post=False)
    ta = hostmirror.upload(a, fs)
    hostmirror.upload(b, fs, tables=ta)
    ca, cb = hostmirror.PlayClocks(fs), hostmirror.PlayClocks(fs)
    fr = hostmirror.scene_frusta(fs)
    wide = fr.copy()
    # one opaque 16-byte record per entry: the survivors come back as their records, gathered on the device
    rec_table = rng.integers(0, 2**32, (a.n_entries[0], 4), dtype=np.uint32)
    with pytest.raises(hgb.HgbError, match="for a table of"):
        a.set_entry_records(0, np.concatenate([rec_table, rec_table[:1]]))
    a.set_entry_records(0, rec_table[:-5])  # a prefix ...
    with pytest.raises(hgb.HgbError, match="no records"):
        a.get_visible_records(0, 0)
    with pytest.raises(hgb.HgbError, match="leave a gap"):
        a.set_entry_records(0, rec_table[-2:], first=len(rec_table) - 2)
    a.set_entry_records(0, rec_table[-5:], first=len(rec_table) - 5)  # ... and the tail behind it
    for frame in range(12):
        if frame == 6:  # a far wider frustum: the visible list grows well past the +25 % guess of the fetch
            wide = hostmirror.make_perspective_frustum(0.1, 5000.0, 2.8, 1.0, 1.0, np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -400], np.float32))
            a.set_views(np.ascontiguousarray(wide, np.float32)); b.set_views(np.ascontiguousarray(wide, np.float32))
        for _ in range(3):  # back to back: tails overlap the following frames
            a.update(ca.step(scenegen.FRAME_DT), fused_cull=True)
            b.update(cb.step(scenegen.FRAME_DT), fused_cull=True)
        ia, ib = a.get_visible(0, 0), b.get_visible(0, 0)
        assert len(ia) == len(ib) and (ia == ib).all(), "frame %d" % frame
        ip = a.get_visible_ptr(0, 0)
        assert len(ip) == len(ia) and (ip == ia).all()
        got = a.get_visible_records(0, 0).view(np.uint32).reshape(-1, 4)
        assert got.shape[0] == len(ia) and (got == rec_table[ia]).all(), "records, frame %d" % frame
        assert (bits(a.get_worlds()) == bits(b.get_worlds())).all()
    assert len(ia) > 0
    a.close(); b.close()


def test_async_tail_chosen_per_frame(hgb):
    """hgb_options.async_post = 0 (default): back-to-back updates run with the asynchronous tail; a caller that reads the result after every
    update gets the one-graph frame from the third such frame on, and the tail comes back with the first update that finds the previous
    result unread. The lists are the same in every mode."""
    rng = np.random.default_rng(94)
    fs = scenegen.static_scene(40_000, 6, 7, n_models=3, seed=94)
    nodes = np.sort(rng.choice(fs.n, 10_000, replace=False)).astype(np.int32)
    scenegen.add_position_tracks(fs, nodes, rng, n_keys=12)
    a, b = hgb.Context(0), hgb.Context(0, async_post=False)
    ta = hostmirror.upload(a, fs)
    hostmirror.upload(b, fs, tables=ta)
    ca, cb = hostmirror.PlayClocks(fs), hostmirror.PlayClocks(fs)

    def frame():
        a.update(ca.step(scenegen.FRAME_DT), fused_cull=True)
        b.update(cb.step(scenegen.FRAME_DT), fused_cull=True)
        return a.frame_was_async

    def same():
        ia, ib = a.get_visible(0, 0), b.get_visible(0, 0)
        assert len(ia) == len(ib) and len(ia) > 0 and (ia == ib).all()

    modes = [frame() for _ in range(4)]  # nobody reads: the tail overlaps the next frame
    assert modes == [True] * 4 and not b.frame_was_async
    same()
    modes = []
    for _ in range(5):  # read after every update
        modes.append(frame())
        same()
    assert modes == [True, False, False, False, False], modes  # (the read before this loop was the first of the streak)
    modes = [frame() for _ in range(3)]  # the first of these still follows a read
    assert modes == [False, True, True], modes
    same()
    a.close(); b.close()


def test_cull_list_rejects_oversized_bounds_table(hgb):
    ctx = hgb.Context(0)
    ctx.set_topology(np.full(4, 0xFFFFFFFF, np.uint32))
    ctx.upload_trs(np.tile(np.array([0, 0, 5, 0, 0, 0, 1, 1, 1], np.float32), (4, 1)))
    ctx.update(None, fused_cull=False)
    fake = np.zeros((1, 6), np.float32)

    class Big:  # claims 2^24 bounds without allocating them: the count alone must be refused
        shape = (1 << 24, 6)
    with pytest.raises(hgb.HgbError, match=r"2\^24-1"):
        n = C.c_uint32(0)
        out = np.zeros(4, np.uint32)
        rc = ctx._L.hgb_cull_list(ctx._h, 1, np.zeros(1, np.uint32).ctypes.data_as(C.c_void_p), np.zeros(1, np.uint32).ctypes.data_as(C.c_void_p), 1 << 24,
                                  fake.ctypes.data_as(C.c_void_p), np.zeros(24, np.float32).ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.byref(n))
        ctx._ck(rc)
    ctx.close()


def test_skinned_list_cull_opt_in(hgb):
    """OPT-IN culling of skinned display lists (the reference never culls them, scene_forward_pipeline.cpp:560-562; neither does the
    default path here). hgb_cull_skinned_list keeps a list unless the union over its bones of bounds[lst] carried through
    world[mtx_idx] * bind_pose[bone] is V_Outside. Expected result: the same definition composed from the oracle port's bit-exact
    operator*(Mat4,Mat4), operator*(Mat4,MinMax) and TestVisibility (min / max of the union are exact)."""
    port = oracle_lib.load_port()
    P = oracle_lib.P
    fs = scenegen.config_c3(n_chars=96, n_bones=64, n_keys=8)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    clocks = hostmirror.PlayClocks(fs)
    for _ in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
    w = ctx.get_worlds()
    n_lists = 96 * 2
    bone_off = (np.arange(n_lists + 1) * 32).astype(np.uint32)
    bounds_idx = np.tile(np.array([0, 1], np.uint32), 96)
    mtx, bind, pose, bounds = t["skin_mtx_slot"], t["skin_bind_idx"], np.ascontiguousarray(t["bind_pose"], np.float32), np.ascontiguousarray(t["bounds"], np.float32)
    assert len(mtx) == n_lists * 32
    got_all = []
    for cam_z in (0.0, -40.0, 150.0):
        fr = hostmirror.make_perspective_frustum(0.1, 90.0, 0.9, 16.0, 9.0, np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, cam_z], np.float32))[0]
        got = ctx.cull_skinned_list(bone_off, mtx, bind, pose, bounds_idx, bounds, fr)
        want = []
        for li in range(n_lists):
            mn, mx = np.full(3, np.inf, np.float32), np.full(3, -np.inf, np.float32)
            for b in range(bone_off[li], bone_off[li + 1]):
                m, box = np.zeros(12, np.float32), np.zeros(6, np.float32)
                port.hgp_mat4_mul(P(np.ascontiguousarray(w[mtx[b]])), P(np.ascontiguousarray(pose[bind[b]])), P(m))
                port.hgp_minmax_transform(P(m), P(np.ascontiguousarray(bounds[bounds_idx[li]])), P(box))
                mn, mx = np.minimum(mn, box[:3]), np.maximum(mx, box[3:])
            if port.hgp_test_visibility(P(np.ascontiguousarray(fr, np.float32)), P(np.concatenate([mn, mx]).astype(np.float32))) != 0:
                want.append(li)
        assert len(got) == len(want) and (got == np.array(want, np.uint32)).all(), cam_z
        got_all.append(len(got))
    assert any(0 < g < n_lists for g in got_all), got_all  # the frusta really cut the crowd
    with pytest.raises(hgb.HgbError, match="at most 32"):
        ctx.cull_skinned_list(np.array([0, 40], np.uint32), mtx[:40], bind[:40], pose, bounds_idx[:1], bounds, fr)
    ctx.close()


def test_incremental_reparent_inside_a_level(hgb, oracles):
    """hgb_set_parents: leaves moved under other parents of the SAME level keep the depth-sorted layout (only parent rows change, no
    rebuild); a move that changes a level falls back to the full rebuild. Worlds, previous worlds and visible lists equal the
    reference's after each kind of edit (Scene::SetTransformParent, component.cpp:184-203)."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    rng = np.random.default_rng(95)
    fs = scenegen.static_scene(5_000, 5, 6, n_models=2, seed=95)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = RefScene(ref, fs)
    fr = hostmirror.scene_frusta(fs)[0]
    depth = np.zeros(fs.n, np.int32)
    for i in range(fs.n):  # parents precede children in static_scene
        depth[i] = depth[fs.parent[i]] + 1 if fs.parent[i] >= 0 else 0
    for _ in range(2):
        ctx.update(None, fused_cull=True); orc.update(0)
    launches_before = ctx.launch_count
    # same-level moves: deepest-level leaves to other parents of the level above
    deepest = np.nonzero(depth == depth.max())[0]
    above = np.nonzero(depth == depth.max() - 1)[0]
    moved = rng.choice(deepest, 150, replace=False).astype(np.uint32)
    new_par = rng.choice(above, 150).astype(np.uint32)
    assert ctx.set_parents(moved, new_par) is False  # no rebuild
    for n_, p_ in zip(moved, new_par):
        ref.hgref_set_parent(orc.h, int(n_), int(p_))
    ctx.update(None, fused_cull=True); orc.update(0)
    assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all()
    assert (bits(ctx.get_worlds(previous=True)) == bits(orc.worlds(previous=True)[: fs.n])).all()
    check_visible(ctx, t, orc, fr)
    assert ctx.launch_count - launches_before < 12  # parent rows + one frame: nothing was rebuilt or re-captured piecemeal
    # a move across levels: rebuild
    up = rng.choice(deepest, 20, replace=False).astype(np.uint32)
    up_par = rng.choice(np.nonzero(depth == 1)[0], 20).astype(np.uint32)
    assert ctx.set_parents(up, up_par) is True
    for n_, p_ in zip(up, up_par):
        ref.hgref_set_parent(orc.h, int(n_), int(p_))
    for _ in range(2):
        ctx.update(None, fused_cull=True); orc.update(0)
    assert (bits(ctx.get_worlds()) == bits(orc.worlds()[: fs.n])).all()
    assert (bits(ctx.get_worlds(previous=True)) == bits(orc.worlds(previous=True)[: fs.n])).all()
    check_visible(ctx, t, orc, fr)
    orc.close(); ctx.close()


def _extend_scene(fs, trs, parent, model):
    """a FlatScene with the given nodes appended (static, single-list opaque objects or none)"""
    import dataclasses
    k = len(parent)
    return dataclasses.replace(
        fs, trs=np.concatenate([fs.trs, trs]).astype(np.float32), parent=np.concatenate([fs.parent, parent]).astype(np.int32),
        disabled=np.concatenate([fs.disabled, np.zeros(k, np.uint8)]), model=np.concatenate([fs.model, model]).astype(np.int32),
        n_mats=np.concatenate([fs.n_mats, np.ones(k, np.uint8)]), transparent_mask=np.concatenate([fs.transparent_mask, np.zeros(k, np.uint32)]))


def test_append_and_remove_nodes_in_place(hgb, oracles):
    """hgb_append_nodes / hgb_append_display_entries / hgb_remove_nodes with append slack: leaves spawned under any node (also under nodes
    spawned a moment ago) and despawned again without a depth sort, a table rebuild or a graph capture; worlds, previous worlds and visible
    lists equal the reference's (CreateNode + SetParent / DestroyNode + GarbageCollect, scene.cpp:92-237) after every step. A chain too long
    for the append region, or no slack at all, falls back to the rebuild with the same results."""
    ref, _ = oracles
    if ref is None:
        pytest.skip("needs the compiled reference")
    from tests.scene_io import RefScene
    rng = np.random.default_rng(97)
    fs = scenegen.static_scene(20_000, 6, 7, n_models=2, seed=97)
    n0 = fs.n
    ctx = hgb.Context(0, append_slack=2048, async_post=True)  # (forced: the launch / rebuild counts below must not see a change of frame mode)
    plain = hgb.Context(0)  # no slack: every edit rebuilds
    t = hostmirror.upload(ctx, fs)
    hostmirror.upload(plain, fs, tables=t)
    orc = RefScene(ref, fs)
    fr = hostmirror.scene_frusta(fs)[0]
    for _ in range(2):
        ctx.update(None, fused_cull=True); plain.update(None, fused_cull=True); orc.update(0)
    check_visible(ctx, t, orc, fr)

    # ---- spawn: 300 leaves under random old nodes, 60 more under freshly spawned ones (chains up to 3 deep), 10 roots
    k = 370
    parent = np.empty(k, np.int32)
    parent[:300] = rng.integers(0, n0, 300)
    parent[300:330] = n0 + rng.integers(0, 300, 30)
    parent[330:360] = n0 + 300 + rng.integers(0, 30, 30)
    parent[360:] = -1
    trs = scenegen.random_trs(rng, k, parent)
    trs[:, 6:9] = rng.uniform(0.8, 1.25, (k, 3))
    trs[360:, 0:3] = rng.uniform(-2, 2, (10, 3))
    model = rng.integers(0, 2, k).astype(np.int32)
    fs2 = _extend_scene(fs, trs, parent, model)
    t2 = hostmirror.build_tables(fs2)
    n_old_ent = len(t["ent0_mtx_slot"])
    assert (t2["ent0_mtx_slot"][:n_old_ent] == t["ent0_mtx_slot"]).all()  # new nodes sit behind the old ones in the display lists
    ref.hgref_add_nodes(orc.h, k, oracle_lib.P(np.ascontiguousarray(trs)), oracle_lib.P(parent), None, oracle_lib.P(model),
                        oracle_lib.P(np.ones(k, np.uint8)), oracle_lib.P(np.zeros(k, np.uint32)), None)
    slots = np.arange(n0, n0 + k, dtype=np.uint32)
    par_slot = np.where(parent >= 0, parent, 0xFFFFFFFF).astype(np.uint32)
    before, rebuilds = ctx.launch_count, ctx.rebuild_count
    assert ctx.append_nodes(slots, par_slot, trs) is False
    assert ctx.append_display_entries(0, t2["ent0_mtx_slot"][n_old_ent:], t2["ent0_bounds_idx"][n_old_ent:], t2["ent0_flags"][n_old_ent:]) is False
    assert plain.append_nodes(slots, par_slot, trs) is True
    assert plain.append_display_entries(0, t2["ent0_mtx_slot"][n_old_ent:], t2["ent0_bounds_idx"][n_old_ent:], t2["ent0_flags"][n_old_ent:]) is True
    for frame in range(2):
        ctx.update(None, fused_cull=True); plain.update(None, fused_cull=True); orc.update(0)
        w_ref = orc.worlds()[: n0 + k]
        assert (bits(ctx.get_worlds()) == bits(w_ref)).all(), "frame %d" % frame
        assert (bits(plain.get_worlds()) == bits(w_ref)).all()
        if frame:  # the first frame after the spawn: the reference's previous rows of new slots are whatever resize() left
            assert (bits(ctx.get_worlds(previous=True)) == bits(orc.worlds(previous=True)[: n0 + k])).all()
        check_visible(ctx, t2, orc, fr)
        check_visible(plain, t2, orc, fr)
    assert ctx.launch_count - before < 16  # two edit kernels + two frames: nothing was re-sorted, rebuilt or captured piecemeal

    # ---- EnableNode / DisableNode of a few entries on live tables: flags only
    tog = rng.choice(len(t2["ent0_flags"]), 40, replace=False)
    for e in tog:
        node = int(t2["ent0_mtx_slot"][e])
        now = 0 if t2["ent0_flags"][e] else 1
        t2["ent0_flags"][e] = now
        ref.hgref_enable_node(orc.h, node, 0 if now else 1)
        ctx.set_entry_flags(0, int(e), np.array([now], np.uint8))
        plain.set_entry_flags(0, int(e), np.array([now], np.uint8))
    ctx.update(None, fused_cull=True); plain.update(None, fused_cull=True); orc.update(0)
    check_visible(ctx, t2, orc, fr)
    check_visible(plain, t2, orc, fr)
    assert ctx.launch_count - before < 16 + 40 + 8
    assert ctx.rebuild_count == rebuilds

    # ---- despawn: childless nodes of the append region and of the sorted part
    has_child = np.zeros(n0 + k, bool)
    has_child[fs2.parent[fs2.parent >= 0]] = True
    gone = np.concatenate([rng.choice(np.nonzero(~has_child[n0:])[0] + n0, 80, replace=False), rng.choice(np.nonzero(~has_child[:n0])[0], 80, replace=False)]).astype(np.uint32)
    for nd in gone:
        ref.hgref_destroy_node(orc.h, int(nd))
    ref.hgref_garbage_collect(orc.h)
    before = ctx.launch_count
    assert ctx.remove_nodes(gone) is False
    assert plain.remove_nodes(gone) is True
    live = np.ones(n0 + k, bool)
    live[gone] = False
    for kind in (0,):
        t2["ent%d_flags" % kind][~live[t2["ent%d_mtx_slot" % kind]]] = 1  # gone from the reference's lists: like a disabled entry for the rank check
    for frame in range(2):
        ctx.update(None, fused_cull=True); plain.update(None, fused_cull=True); orc.update(0)
        w_ref = orc.worlds()[: n0 + k]
        assert (bits(ctx.get_worlds())[live] == bits(w_ref)[live]).all()
        assert (bits(plain.get_worlds())[live] == bits(w_ref)[live]).all()
        check_visible(ctx, t2, orc, fr)
        check_visible(plain, t2, orc, fr)
    assert ctx.launch_count - before < 12 and ctx.rebuild_count == rebuilds
    # the stand-alone cull reads the entry tables: the removed nodes' entries are disabled there as well
    ctx.update(None, fused_cull=False); ctx.cull(); orc.update(0)
    check_visible(ctx, t2, orc, fr)

    # ---- freed positions of the append region are handed out again; a chain too long for the region rebuilds
    back = gone[:40]  # 40 of the despawned slots of the append region
    ptrs = scenegen.random_trs(rng, 40, np.zeros(40, np.int32))
    ppar = rng.integers(0, n0, 40).astype(np.uint32)
    ppar[np.isin(ppar, gone)] = 0
    assert ctx.append_nodes(back, ppar, ptrs) is False
    # a slot freed in the SORTED part is taken again in place when the new parent sits in an earlier level (here: a root) ...
    depth0 = np.zeros(n0, np.int32)
    for i in range(n0):
        depth0[i] = depth0[fs.parent[i]] + 1 if fs.parent[i] >= 0 else 0
    srt = gone[80:100]  # despawned slots of the sorted part (childless nodes: deeper than level 0)
    assert (depth0[srt] > 0).all()
    strs = scenegen.random_trs(rng, 20, np.zeros(20, np.int32))
    roots0 = np.nonzero(fs.parent < 0)[0]
    spar = rng.choice(roots0, 20).astype(np.uint32)
    assert ctx.append_nodes(srt, spar, strs) is False
    back = np.concatenate([back, srt]); ptrs = np.concatenate([ptrs, strs]); ppar = np.concatenate([ppar, spar])
    chain = np.arange(n0 + k, n0 + k + 14, dtype=np.uint32)
    cpar = np.concatenate([[5], chain[:-1]]).astype(np.uint32)
    ctrs = scenegen.random_trs(rng, 14, np.zeros(14, np.int32))
    ctrs[:, 6:9] = 1.0
    assert ctx.append_nodes(chain, cpar, ctrs) is True  # 13 ancestors inside the region > kMaxWalk - 1
    # the same final hierarchy built from scratch
    fs3 = _extend_scene(fs2, ctrs, cpar.astype(np.int32), np.full(14, -1, np.int32))
    fs3.trs[back] = ptrs
    fs3.parent[back] = ppar.astype(np.int32)
    scratch = hgb.Context(0)
    par3 = np.where(fs3.parent >= 0, fs3.parent, 0xFFFFFFFF).astype(np.uint32)
    used3 = np.ones(fs3.n, np.uint8)
    used3[gone[40:80]] = 0
    used3[gone[100:]] = 0
    scratch.set_topology(par3, used3)
    scratch.upload_trs(np.ascontiguousarray(fs3.trs, np.float32))
    for _ in range(2):
        ctx.update(None, fused_cull=False); scratch.update(None, fused_cull=False)
    sel = used3.astype(bool)
    assert (bits(ctx.get_worlds())[sel] == bits(scratch.get_worlds())[sel]).all()
    orc.close(); ctx.close(); plain.close(); scratch.close()
