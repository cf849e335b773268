"""GPU parity tests: the CUDA path, called through the C ABI (libhgb200.so), against the oracle on the same seeded scenes.

Oracle = the UNMODIFIED reference compiled into oracle/_ref/libhg_ref.so when present (it travels to the GPU box prebuilt), else the
plain-C port, which is pinned bit-for-bit against it (tests/test_oracle_port_vs_ref.py, tests/golden).
Bars: visible lists / counts bit-exact; world and bone matrices within 1e-5 relative (north_star) — and in fact bit-identical for
every path that does not go through libm's double atan2 (quaternion rotation tracks), which is asserted separately.
"""
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
