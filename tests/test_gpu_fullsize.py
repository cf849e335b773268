"""GPU parity at BASELINE.json's FULL sizes (configs[2..4]). The oracle cannot run 16M nodes in seconds, so each test combines
  (1) a bit-exact comparison against the oracle on a complete, independent part of the batch (one scene of the c4 batch / a block of
      c5 instances / a block of c3 characters — scenes of a batch share nothing, so the part is a stand-alone reference scene), and
  (2) size-independent properties over the WHOLE batch: the visible list is strictly ascending (stable compaction), its length equals
      the count word, disabled entries never survive, per-instance offsets are monotone and end at the count, the stand-alone cull
      (hgb_cull) reproduces the fused pass exactly (idempotence), previous-frame worlds equal last frame's worlds, and both level
      kernel sets (default: one thread per node, one launch per level, sampler first / A-B: persistent TMA-staged levels, merged small
      levels, sampler overlapped by target level on a second stream over 1M-node slices) give identical bits.
"""
import numpy as np
import pytest

from harfang3d_b200 import hostmirror, scenegen
from tests import oracle_lib
from tests.oracle_lib import bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hgb():
    import harfang3d_b200
    return harfang3d_b200


def _oracle(fs):
    ref = oracle_lib.load_ref()
    if ref is not None:
        from tests.scene_io import RefScene
        return RefScene(ref, fs)
    from tests.port_scene import PortScene
    return PortScene(oracle_lib.load_port(), fs)


def _list_properties(ctx, t, kind=0, view=0):
    idx = ctx.get_visible(kind, view)
    assert ctx.get_visible_count(kind, view) == len(idx)
    assert (np.diff(idx.astype(np.int64)) > 0).all(), "visible list must be strictly ascending"
    assert (t["ent%d_flags" % kind][idx] == 0).all(), "a disabled entry survived"
    return idx


def test_c4_full_16m_batch(hgb):
    """configs[3]: 16 scenes x 1M nodes, depth 8, 25 % animated. Scene 5 of the batch is checked bit-for-bit against the oracle over 3
    frames; the whole batch through properties."""
    n_per = 1_000_000
    fs = scenegen.config_c4(seed=4, n_scenes=16, nodes_per_scene=n_per, animated_frac=0.25, n_keys=16)
    k = 5
    sub = scenegen.extract_range(fs, k * n_per, (k + 1) * n_per)
    orc = _oracle(sub)
    ctx = hgb.Context(0)
    ctx2 = hgb.Context(0, tma_level=True, top_walk=-1, overlap_sampler=True, slice_nodes=1 << 20)
    t = hostmirror.upload(ctx, fs)
    hostmirror.upload(ctx2, fs, tables=t)
    clocks, clocks2 = hostmirror.PlayClocks(fs), hostmirror.PlayClocks(fs)
    fr = hostmirror.scene_frusta(fs)[0]
    ent_node = t["ent0_mtx_slot"]
    prev = None
    for frame in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
        ctx2.update(clocks2.step(scenegen.FRAME_DT), fused_cull=True)
        orc.update(scenegen.FRAME_DT)
        w = ctx.get_worlds(k * n_per, n_per)
        assert (bits(w) == bits(orc.worlds()[:n_per])).all(), "frame %d: worlds of scene %d differ from the oracle" % (frame, k)
        idx = _list_properties(ctx, t)
        _, mtx_ref = orc.cull(0, fr)
        mine = ent_node[idx]
        mine = mine[(mine >= k * n_per) & (mine < (k + 1) * n_per)] - k * n_per
        assert len(mine) == len(mtx_ref) and (mine == mtx_ref).all(), "frame %d: visible list of scene %d differs" % (frame, k)
        assert 0 < len(idx) < len(ent_node)
        # both level kernels: identical bits over the whole batch
        assert (ctx2.get_visible(0, 0) == idx).all()
        w_all = ctx.get_worlds()
        assert (bits(ctx2.get_worlds()) == bits(w_all)).all()
        if prev is not None:
            assert (bits(ctx.get_worlds(previous=True)) == bits(prev)).all()
        prev = w_all
        # idempotence: the stand-alone cull over the stored worlds gives the same list
        ctx.cull()
        assert (ctx.get_visible(0, 0) == idx).all()
    orc.close(); ctx.close(); ctx2.close()


def test_c5_full_instance_batch(hgb):
    """configs[4] per-GPU share at 2 GPUs: 4096 instances x 2048 nodes, one camera per instance. Instances 1000..1063 are checked
    against the oracle (one reference scene holding the 64 instances); the whole batch through properties."""
    fs = scenegen.config_c5(seed=5, n_instances=4096)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    ctx.update(None, fused_cull=True)
    idx = _list_properties(ctx, t)
    off = ctx.get_instance_offsets().astype(np.int64)
    assert off[0] == 0 and off[-1] == len(idx) and (np.diff(off) >= 0).all()
    ent_node = t["ent0_mtx_slot"]
    inst_of_entry = np.searchsorted(fs.instance_start, ent_node[idx], side="right") - 1
    assert (np.diff(inst_of_entry) >= 0).all()
    assert (np.bincount(inst_of_entry, minlength=4096) == np.diff(off)).all(), "per-instance offsets do not match the list"
    i0, i1 = 1000, 1064
    sub = scenegen.slice_instances(fs, i0, i1)
    orc = _oracle(sub)
    orc.update(0)
    lo, hi = int(fs.instance_start[i0]), int(fs.instance_start[i1])
    assert (bits(ctx.get_worlds(lo, hi - lo)) == bits(orc.worlds()[: hi - lo])).all()
    frusta = hostmirror.scene_frusta(sub)
    for i in range(i1 - i0):
        _, mtx_ref = orc.cull(0, frusta[i])
        a, b = sub.instance_start[i], sub.instance_start[i + 1]
        want = mtx_ref[(mtx_ref >= a) & (mtx_ref < b)] + lo
        got = ent_node[idx[off[i0 + i]:off[i0 + i + 1]]]
        assert len(got) == len(want) and (got == want).all(), "instance %d" % (i0 + i)
    ctx.cull()
    assert (ctx.get_visible(0, 0) == idx).all() and (ctx.get_instance_offsets() == off).all()
    orc.close(); ctx.close()


def test_c3_full_4096_characters(hgb):
    """configs[2]: 4096 skinned characters x 64 bones, Position/Rotation/Scale tracks per bone (half of the characters with quaternion
    rotation tracks), bone matrices with the previous-frame twin. The whole scene runs in the oracle for 3 frames (270k nodes)."""
    fs = scenegen.config_c3(seed=3, n_chars=4096, n_bones=64, n_keys=32)
    ctx = hgb.Context(0)
    t = hostmirror.upload(ctx, fs)
    orc = _oracle(fs)
    clocks = hostmirror.PlayClocks(fs)
    fr = hostmirror.scene_frusta(fs)[0]
    per = 66
    euler_chars = np.arange(0, 4096, 2)
    euler_nodes = (euler_chars[:, None] * per + np.arange(per)[None, :]).reshape(-1)
    for frame in range(3):
        ctx.update(clocks.step(scenegen.FRAME_DT), fused_cull=True)
        ctx.bone_matrices(with_previous=True)
        orc.update(scenegen.FRAME_DT)
        w, w_ref = ctx.get_worlds(), orc.worlds()[: fs.n]
        # bit-identical, Euler AND quaternion characters (the quaternion path's double atan2 is narrowed exactly, hgb_atan2_exact.h)
        assert (bits(w) == bits(w_ref)).all(), "frame %d: %d world rows differ" % (frame, int((bits(w) != bits(w_ref)).any(axis=1).sum()))
        b, b_ref = ctx.get_bone_matrices(), orc.bone_matrices()
        assert b.shape == b_ref.shape == (4096 * 64, 16)
        assert (bits(b) == bits(np.ascontiguousarray(b_ref, np.float32))).all()
        if frame:
            bp, bp_ref = ctx.get_bone_matrices(previous=True), orc.bone_matrices(previous=True)
            assert (bits(bp) == bits(np.ascontiguousarray(bp_ref, np.float32))).all()
        idx = _list_properties(ctx, t)
        _, mtx_ref = orc.cull(0, fr)
        got = t["ent0_mtx_slot"][idx]
        assert len(got) == len(mtx_ref) and (got == mtx_ref).all(), "frame %d: visible list differs from the oracle" % frame
    orc.close(); ctx.close()
