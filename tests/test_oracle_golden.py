"""CPU tests (no GPU, no /root/reference): the oracle port (oracle/hg_port.c) against
  (1) golden vectors generated from the UNMODIFIED reference (tests/golden/*.npz, made by tests/golden/gen_golden.py), bit-exact;
  (2) the known-answer checks of the reference's own unit tests for this path, restated with the reference's
      file:line and tolerance.
"""
import ctypes as C
import os

import numpy as np
import pytest

from harfang3d_b200 import hostmirror
from tests.golden.gen_golden import golden_scenes, scene_digest
from tests.oracle_lib import P, bits
from tests.port_scene import PortScene

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gm():
    return np.load(os.path.join(G, "math.npz"))


@pytest.fixture(scope="module")
def gs():
    return np.load(os.path.join(G, "scenes.npz"))


# ---- (1) golden vectors from the reference ----------------------------------------------------------------------------------
def test_golden_sincos(port, gm):
    s = np.array([port.hgp_sinf(float(x)) for x in gm["sincos_x"]], np.float32)
    c = np.array([port.hgp_cosf(float(x)) for x in gm["sincos_x"]], np.float32)
    assert (bits(s) == bits(gm["sinf"])).all() and (bits(c) == bits(gm["cosf"])).all()


def test_golden_matrix_functions(port, gm):
    trs, n = gm["trs"], len(gm["trs"])
    m, o = np.zeros((n, 12), np.float32), np.zeros(12, np.float32)
    for i in range(n):
        port.hgp_transformation_mat4(P(trs[i]), P(m[i]))
    assert (bits(m) == bits(gm["trs_mat4"])).all()
    for i in range(n):
        port.hgp_mat4_mul(P(m[i]), P(m[(i * 7 + 1) % n]), P(o))
        assert (bits(o) == bits(gm["mat4_mul"][i])).all()
        port.hgp_inverse_fast(P(m[i]), P(o))
        assert (bits(o) == bits(gm["inverse_fast"][i])).all()


def test_golden_minmax_frustum_visibility(port, gm):
    n = len(gm["trs"])
    o6, fr, cam = np.zeros(6, np.float32), np.zeros(24, np.float32), np.zeros(12, np.float32)
    seen = set()
    for i in range(n):
        port.hgp_minmax_transform(P(np.ascontiguousarray(gm["mat4_mul"][i])), P(np.ascontiguousarray(gm["minmax_in"][i])), P(o6))
        assert (bits(o6) == bits(gm["minmax_out"][i])).all()
        port.hgp_transformation_mat4(P(np.ascontiguousarray(gm["frustum_cam_trs"][(i * 3) % n])), P(cam))
        port.hgp_make_perspective_frustum(*[np.float32(x) for x in gm["frustum_args"][i]], P(cam), P(fr))
        assert (bits(fr) == bits(gm["frustum"][i])).all()
        v = port.hgp_test_visibility(P(fr), P(o6))
        assert v == gm["visibility"][i]
        seen.add(v)
    assert seen == {0, 1, 2}
    # the vectorised host-side frustum builder used by the product's host mirror gives the same planes
    cams = np.zeros((n, 12), np.float32)
    for i in range(n):
        port.hgp_transformation_mat4(P(np.ascontiguousarray(gm["frustum_cam_trs"][(i * 3) % n])), P(cams[i]))
    a = gm["frustum_args"]
    for i in range(0, n, 50):
        got = hostmirror.make_perspective_frustum(a[i, 0], a[i, 1], a[i, 2], a[i, 3], a[i, 4], cams[i])
        assert (bits(got[0]) == bits(gm["frustum"][i])).all()


def test_golden_quat_to_euler(port, gm):
    o = np.zeros(3, np.float32)
    for q, e in zip(gm["quat"], gm["quat_euler"]):
        port.hgp_quat_to_euler(P(np.ascontiguousarray(q)), P(o))
        assert (bits(o) == bits(e)).all()


def test_golden_tracks(port, gm):
    for i in range(len(gm["tk_n"])):
        nk = int(gm["tk_n"][i])
        t, v, ten, bia, q = (np.ascontiguousarray(gm[k][i]) for k in ("tk_t", "tk_v", "tk_ten", "tk_bias", "tk_q"))
        for j, at in enumerate(gm["tk_at"][i]):
            o3, oq = np.full(3, 7, np.float32), np.zeros(4, np.float32)
            r = port.hgp_eval_vec3_track(nk, P(t), P(v), P(ten), P(bia), int(at), P(o3))
            assert r == gm["tk_ret"][i][j] and (bits(o3) == bits(gm["tk_out3"][i][j])).all()
            k0, k1 = C.c_int(-1), C.c_int(-1)
            inside = port.hgp_get_interval_keys(nk, P(t), int(at), C.byref(k0), C.byref(k1))
            kf = gm["tk_kf"][i][j]
            assert inside == kf[2] and k0.value == kf[0] and (not inside or k1.value == kf[1])
            port.hgp_eval_quat_track(nk, P(t), P(q), int(at), P(oq))
            assert (bits(oq) == bits(gm["tk_outq"][i][j])).all()


@pytest.mark.parametrize("name", list(golden_scenes()))
def test_golden_scenes_port(port, gs, name):
    """whole frames: host mirror (tables, clocks, frusta) + port arithmetic == reference Scene::Update / GetModelDisplayLists / Cull"""
    make, frames, dt = golden_scenes()[name]
    fs = make()
    assert scene_digest(fs) == gs[name + "_digest"], "scene generator drifted: regenerate tests/golden with gen_golden.py"
    sc = PortScene(port, fs)
    for _ in range(frames):
        sc.update(dt)
    assert (bits(sc.worlds()) == bits(gs[name + "_worlds"])).all()
    if frames > 1:
        assert (bits(sc.worlds(previous=True)) == bits(gs[name + "_worlds_prev"])).all()
    if name + "_trs" in gs:
        assert (bits(sc.trs(gs[name + "_trs_nodes"])) == bits(gs[name + "_trs"])).all()
    # display-list table of the host mirror == GetModelDisplayLists (enabled entries, emission order)
    t = sc.t
    assert (t["ent0_mtx_slot"][t["ent0_flags"] == 0] == gs[name + "_dl_mtx"]).all()
    assert gs[name + "_dl_counts"][0] == int((t["ent0_flags"] == 0).sum())
    frusta = hostmirror.scene_frusta(fs)
    for v in range(len(frusta)):
        pos, mtx = sc.cull(0, frusta[v])
        assert (pos == gs["%s_vis%d_pos" % (name, v)]).all() and (mtx == gs["%s_vis%d_mtx" % (name, v)]).all()
    if name + "_bones" in gs:
        assert gs[name + "_dl_counts"][2] * 32 == len(t["skin_mtx_slot"])
        assert (bits(sc.bone_matrices()) == bits(gs[name + "_bones"])).all()
        assert (bits(sc.bone_matrices(previous=True)) == bits(gs[name + "_bones_prev"])).all()


def test_plane_margin_statistic(port):
    """hostmirror.plane_margins (SURVEY.md §8d diagnostic): a float64 re-evaluation of the cull predicate classifies every entry whose margin
    is above float32 noise exactly like the oracle's visible list, and the reported minimum is the smallest of those margins"""
    from harfang3d_b200 import scenegen
    fs = scenegen.config_c1()
    sc = PortScene(port, fs)
    sc.update(0)
    fr = hostmirror.scene_frusta(fs)[0]
    m_out, m_clip, n = hostmirror.plane_margins(sc.t, sc.worlds(), fr)
    assert n == int((sc.t["ent0_flags"] == 0).sum()) and 0 <= m_out < 1 and 0 <= m_clip < 1  # 10k boxes against 6 planes: some come close
    _, _, n4 = hostmirror.plane_margins(sc.t, sc.worlds(), fr, sample=4)
    assert n4 == (n + 3) // 4
    # independent float64 classification of the enabled entries
    t = sc.t
    sel = np.nonzero(t["ent0_flags"] == 0)[0]
    W = sc.worlds().reshape(-1, 3, 4)[t["ent0_mtx_slot"][sel]].astype(np.float64)
    b = t["bounds"][t["ent0_bounds_idx"][sel]].astype(np.float64)
    lo, hi = W[:, :, :3] * b[:, None, :3], W[:, :, :3] * b[:, None, 3:]
    mn, mx = W[:, :, 3] + np.minimum(lo, hi).sum(2), W[:, :, 3] + np.maximum(lo, hi).sum(2)
    f = fr.reshape(6, 4).astype(np.float64)
    s_out = (mn + mx) @ f[:, :3].T - (mx - mn) @ np.abs(f[:, :3]).T + 2 * f[:, 3]
    assert np.abs(s_out).min() == pytest.approx(m_out, rel=1e-12)
    inside64 = ~(s_out > 0).any(1)
    pos, _ = sc.cull(0, fr)
    inside_port = np.zeros(len(sel), bool)
    inside_port[pos] = True
    sure = np.abs(s_out).min(1) > 1e-3
    assert sure.sum() > 0.99 * len(sel) and (inside64[sure] == inside_port[sure]).all()


# ---- (2) the reference's own unit-test expectations ---------------------------------------------------------------------------
def almost(a, b, eps):
    return np.all(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)) <= eps)


def test_ref_unit_mat4_mul_golden_columns(port):
    """harfang/tests/foundation/mat4.cpp:127-134 — Mat4(row-major ctor) product, columns at 1e-5.
    The reference ctor takes (m00,m10,m20, m01,m11,m21, ...) column by column (matrix4.h)."""
    def from_cols(v):
        m = np.zeros((3, 4), np.float32)
        for c in range(4):
            m[:, c] = v[c * 3:c * 3 + 3]
        return m.reshape(12)
    m = from_cols(np.arange(1, 13, dtype=np.float32))
    n = from_cols((np.arange(1, 13) / 10).astype(np.float32))
    p = np.zeros(12, np.float32)
    port.hgp_mat4_mul(P(m), P(n), P(p))
    p = p.reshape(3, 4)
    assert almost(p[:, 0], [3, 3.6, 4.2], 1e-5) and almost(p[:, 1], [6.6, 8.1, 9.6], 1e-5)
    assert almost(p[:, 2], [10.2, 12.6, 15], 1e-5) and almost(p[:, 3], [23.8, 28.1, 32.4], 1e-5)


def test_ref_unit_inverse_fast(port):
    """harfang/tests/foundation/mat4.cpp:203-209 — m * InverseFast(m) == identity at 1e-5"""
    u = np.float32(2) * np.sqrt(np.float32(2))
    cols = np.array([u, 0, -u, 2, u, 2, 2, -u, 2, -1, 0.5, 2], np.float32)
    m = np.zeros((3, 4), np.float32)
    for c in range(4):
        m[:, c] = cols[c * 3:c * 3 + 3]
    m = m.reshape(12)
    inv, n = np.zeros(12, np.float32), np.zeros(12, np.float32)
    port.hgp_inverse_fast(P(m), P(inv))
    port.hgp_mat4_mul(P(m), P(inv), P(n))
    assert almost(n.reshape(3, 4), np.eye(3, 4), 1e-5)


def test_ref_unit_minmax_transform_vs_brute_force(port):
    """harfang/tests/foundation/minmax.cpp:127-151 — trs * MinMax equals the 8-corner brute force at 1e-6"""
    deg = np.float32(np.pi) / np.float32(180)
    trs = np.array([-0.5, 0.5, 0.8, 45 * deg, -30 * deg, 60 * deg, 2, 0.8, 1], np.float32)
    m, o = np.zeros(12, np.float32), np.zeros(6, np.float32)
    port.hgp_transformation_mat4(P(trs), P(m))
    port.hgp_minmax_transform(P(m), P(np.array([-1, -1, -1, 1, 1, 1], np.float32)), P(o))
    M = m.reshape(3, 4).astype(np.float64)
    corners = np.array([[x, y, z, 1] for x in (-1, 1) for y in (-1, 1) for z in (-1, 1)], np.float64) @ M.T
    assert almost(o[:3], corners.min(0), 1e-6 * 4) and almost(o[3:], corners.max(0), 1e-6 * 4)


def test_ref_unit_frustum_visibility_table(port):
    """harfang/tests/foundation/frustum.cpp:17-54 — 4 boxes x 2 frusta: Inside/Clipped/Clipped/Outside and Inside/Outside/Clipped/Clipped"""
    deg = np.float32(np.pi) / np.float32(180)
    ident = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
    cam1 = np.zeros(12, np.float32)
    port.hgp_transformation_mat4(P(np.array([0, 5, 5, 90 * deg, 0, 0, 1, 1, 1], np.float32)), P(cam1))  # frustum2 of the test (== frustum1 at 1e-5)
    f0, f1 = np.zeros(24, np.float32), np.zeros(24, np.float32)
    args = (np.float32(0.1), np.float32(10), np.float32(60 * deg), np.float32(2560 / 1440), np.float32(1))
    port.hgp_make_perspective_frustum(*args, P(ident), P(f0))
    port.hgp_make_perspective_frustum(*args, P(cam1), P(f1))
    boxes = np.array([[-0.1, -0.1, 4.9, 0.1, 0.1, 5.1], [-0.1, -0.1, 9.9, 0.1, 0.1, 10.1], [-2, -2, 2, 2, 2, 6], [2, -5.1, 4.9, 2.2, -4.8, 5.1]], np.float32)
    OUT, IN, CLIP = 0, 1, 2
    assert [port.hgp_test_visibility(P(f0), P(b)) for b in boxes] == [IN, CLIP, CLIP, OUT]
    assert [port.hgp_test_visibility(P(f1), P(b)) for b in boxes] == [IN, OUT, CLIP, CLIP]


def test_ref_unit_interval_keys(port):
    """harfang/tests/engine/animation.cpp:51-64 — keys at 200/400/600 ms (the 800 ms key was deleted at :46)"""
    ms = 1_000_000
    t = np.array([200, 400, 600], np.int64) * ms
    k0, k1 = C.c_int(-1), C.c_int(-1)
    assert port.hgp_get_interval_keys(3, P(t), 600 * ms, C.byref(k0), C.byref(k1)) == 0 and k0.value == 2  # past end (t[1]+200ms == last key)
    assert port.hgp_get_interval_keys(3, P(t), 500 * ms, C.byref(k0), C.byref(k1)) == 1 and (k0.value, k1.value) == (1, 2)
    assert port.hgp_get_interval_keys(3, P(t), 2000 * ms, C.byref(k0), C.byref(k1)) == 0 and k0.value == 2
    assert port.hgp_get_interval_keys(3, P(t), 20 * ms, C.byref(k0), C.byref(k1)) == 0 and k0.value == 0


def test_ref_unit_vec3_track_at_and_outside_keys(port):
    """harfang/tests/engine/animation.cpp:191-243 — exact key values at key times, clamped outside, false on an empty track"""
    ms = 1_000_000
    t = np.array([200, 400, 600, 800], np.int64) * ms
    v = np.array([[-1, -1, -1], [-1, 0, -1], [0, 1, 0], [0, 1, 0]], np.float32)
    ten, bia = np.array([0.5, 0.75, 0.25, 0], np.float32), np.array([0.5, 0.25, 0.75, 0], np.float32)
    o = np.zeros(3, np.float32)
    assert port.hgp_eval_vec3_track(0, None, None, None, None, 600 * ms, P(o)) == 0
    for at, want in [(100, v[0]), (200, v[0]), (400, v[1]), (600, v[2]), (800, v[3]), (1000, v[3])]:
        assert port.hgp_eval_vec3_track(4, P(t), P(v), P(ten), P(bia), at * ms, P(o)) == 1
        assert (o == want).all(), at


def test_ref_unit_display_list_counts_under_enable_disable(port):
    """harfang/tests/engine/scene.cpp:115-154 — 6 object nodes: 6 / 3 / 6 / 0 opaque lists as nodes are disabled and enabled.
    The host mirror keeps disabled nodes in the entry table with HGB_ENTRY_DISABLED; the count is the number of unflagged entries."""
    from harfang3d_b200 import scenegen
    fs = scenegen.static_scene(6, 1, 1, n_models=1, seed=0, disabled_frac=0.0) if False else None
    n = 6
    fs = scenegen.FlatScene(trs=np.tile(np.array([0, 0, 0, 0, 0, 0, 1, 1, 1], np.float32), (n, 1)), parent=np.full(n, -1, np.int32),
                            disabled=np.zeros(n, np.uint8), model=np.zeros(n, np.int32), n_mats=np.ones(n, np.uint8),
                            transparent_mask=np.zeros(n, np.uint32), models=scenegen.box_models(np.random.default_rng(0), 1),
                            camera=dict(scenegen.DEFAULT_CAMERA))
    def count():
        return int((hostmirror.build_tables(fs)["ent0_flags"] == 0).sum())
    assert count() == 6
    fs.disabled[[0, 2, 4]] = 1
    assert count() == 3
    fs.disabled[:] = 0
    assert count() == 6
    fs.disabled[:] = 1
    assert count() == 0
