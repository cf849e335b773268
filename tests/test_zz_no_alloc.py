"""hgb_compact_visible_noalloc — the allocation-free compaction (static __device__ scratch, csrc/hgb_noalloc.cu) — against the existing
entry points (hgb_update with the fused cull + hgb_get_visible), through the C ABI, on calls in a row that share the scratch.

Collected last on purpose (file name): a failure here hides nothing behind it. The test allocates
its own device buffers with torch; the entry point under test allocates nothing."""
import numpy as np
import pytest

from harfang3d_b200 import capi, hostmirror, scenegen


def _frusta(fs):
    """three different views of the same scene: the camera's own frustum, one moved back, one looking from the side"""
    cam = fs.camera
    views = [hostmirror.scene_frusta(fs)[0]]
    for world12 in (np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -60], np.float32), np.array([0, 0, 1, -30, 0, 1, 0, 0, -1, 0, 0, 0], np.float32)):
        views.append(hostmirror.make_perspective_frustum(cam["znear"], cam["zfar"], cam["fov"], cam["ar_x"], cam["ar_y"], world12))
    return np.ascontiguousarray(np.stack([np.asarray(v, np.float32).reshape(24) for v in views]))


@pytest.mark.gpu
def test_compact_visible_noalloc_matches_the_context_path():
    import torch
    import harfang3d_b200 as hgb

    fs = scenegen.static_scene(100_000, 6, 7, n_models=3, seed=171)
    ctx = hgb.Context(0)
    hostmirror.upload(ctx, fs)
    ctx.set_views(_frusta(fs))
    for _ in range(2):
        ctx.update(None, fused_cull=True)
    n = ctx.n_entries[0]
    assert n > 3 * 8192 and n % 8192 != 0  # several tiles and a ragged tail
    want = [ctx.get_visible(0, v).copy() for v in range(3)]  # the existing entry points' lists, one per view
    assert all(len(w) > 0 for w in want) and len({len(w) for w in want}) > 1
    vis = np.zeros(n, np.uint8)  # the visibility bytes those lists stand for: bit v = survives view v
    for v, w in enumerate(want):
        vis[w] |= np.uint8(1 << v)
    ctx.close()

    dev = torch.device("cuda:0")
    d_vis = torch.from_numpy(vis).to(dev)
    d_out = torch.full((n,), -1, dtype=torch.int32, device=dev)
    d_count = torch.zeros(1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def run(n_entries, bit, capacity):
        d_out.fill_(-1)
        d_count.fill_(-1)
        capi.compact_visible_noalloc(0, d_vis.data_ptr(), n_entries, bit, d_out.data_ptr(), capacity, d_count.data_ptr(), stream)
        torch.cuda.synchronize(dev)
        return d_out.cpu().numpy().view(np.uint32), int(d_count.item())

    # calls in a row on one stream: every one starts from the previous call's ticket and status words
    for rep in range(2):
        for v in range(3):
            out, cnt = run(n, v, n)
            assert cnt == len(want[v]), "view %d, pass %d" % (v, rep)
            assert (out[:cnt] == want[v]).all()  # integer work: bit-exact, same order (ascending = the reference's remove_if order)
            assert (out[cnt:] == 0xFFFFFFFF).all()  # nothing written past the list
    # a shorter table right after a longer one (stale status words of the tiles beyond it), and a prefix that ends inside a tile
    for n_part in (n // 3, 8192, 8191, 33, 1):
        out, cnt = run(n_part, 1, n)
        ref = want[1][want[1] < n_part]
        assert cnt == len(ref) and (out[:cnt] == ref).all()
    # writes stop at the capacity, the count still reports every survivor
    cap = len(want[0]) // 2
    out, cnt = run(n, 0, cap)
    assert cnt == len(want[0]) and (out[:cap] == want[0][:cap]).all() and (out[cap:] == 0xFFFFFFFF).all()
    # an empty table
    out, cnt = run(0, 0, n)
    assert cnt == 0 and (out == 0xFFFFFFFF).all()
    # random bytes, all eight bits, against numpy
    rng = np.random.default_rng(5)
    vis2 = rng.integers(0, 256, n, dtype=np.uint8)
    d_vis.copy_(torch.from_numpy(vis2))
    for bit in (0, 7):
        out, cnt = run(n, bit, n)
        ref = np.nonzero((vis2 >> bit) & 1)[0].astype(np.uint32)
        assert cnt == len(ref) and (out[:cnt] == ref).all()

    # refusals: beyond the static scratch's maximum, bad bit, misaligned input — all before anything is touched
    with pytest.raises(capi.HgbError, match="exceed the maximum"):
        capi.compact_visible_noalloc(0, d_vis.data_ptr(), capi.NOALLOC_MAX_ENTRIES + 1, 0, d_out.data_ptr(), n, d_count.data_ptr(), stream)
    with pytest.raises(capi.HgbError, match="view bit"):
        capi.compact_visible_noalloc(0, d_vis.data_ptr(), n, 8, d_out.data_ptr(), n, d_count.data_ptr(), stream)
    with pytest.raises(capi.HgbError, match="aligned"):
        capi.compact_visible_noalloc(0, d_vis.data_ptr() + 1, n - 1, 0, d_out.data_ptr(), n, d_count.data_ptr(), stream)
    # ... and the call after the refusals still works
    out, cnt = run(n, 7, n)
    assert cnt == int(((vis2 >> 7) & 1).sum())


def test_compact_visible_noalloc_refusals_need_no_device():
    """the size / argument checks come before any CUDA call, and without a device the entry point fails loudly (no CPU path)"""
    buf = np.zeros(64, np.uint8)
    out = np.zeros(64, np.uint32)
    cnt = np.zeros(1, np.uint32)
    aligned = (buf.ctypes.data + 15) & ~15
    with pytest.raises(capi.HgbError, match="exceed the maximum"):
        capi.compact_visible_noalloc(0, aligned, capi.NOALLOC_MAX_ENTRIES + 1, 0, out.ctypes.data, 64, cnt.ctypes.data)
    with pytest.raises(capi.HgbError, match="view bit"):
        capi.compact_visible_noalloc(0, aligned, 16, 9, out.ctypes.data, 64, cnt.ctypes.data)
    with pytest.raises(capi.HgbError, match="null buffer"):
        capi.compact_visible_noalloc(0, aligned, 16, 0, out.ctypes.data, 64, 0)
    with pytest.raises(capi.HgbError, match="aligned"):
        capi.compact_visible_noalloc(0, aligned + 1, 16, 0, out.ctypes.data, 64, cnt.ctypes.data)
    if capi.lib().hgb_device_count() == 0:  # this container: nothing may be computed on the host
        with pytest.raises(capi.HgbError, match="no CPU path"):
            capi.compact_visible_noalloc(0, aligned, 16, 0, out.ctypes.data, 64, cnt.ctypes.data)
        assert (out == 0).all() and cnt[0] == 0
