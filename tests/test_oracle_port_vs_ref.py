"""Pins the plain-C port (oracle/hg_port.c) bit-for-bit against the UNMODIFIED reference (oracle/_ref).
Skipped where the reference could not be built (no /root/reference)."""
import ctypes as C

import numpy as np
import pytest

from tests.oracle_lib import P, bits

pytestmark = pytest.mark.ref


def rand_trs(rng, n):
    t = np.empty((n, 9), np.float32)
    t[:, 0:3] = rng.uniform(-50, 50, (n, 3))
    t[:, 3:6] = rng.uniform(-np.pi, np.pi, (n, 3))
    t[:, 6:9] = rng.uniform(0.5, 2, (n, 3))
    return t


def test_sincos_sampled(port, ref):
    rng = np.random.default_rng(1)
    xs = np.concatenate([
        rng.uniform(-np.pi, np.pi, 200000), rng.uniform(-200, 200, 50000), rng.normal(0, 1e-3, 10000),
        (rng.integers(0, 2**32, 200000, dtype=np.uint64).astype(np.uint32)).view(np.float32).astype(np.float64),
    ]).astype(np.float32)
    xs = xs[np.isfinite(xs)]
    for x in xs[::7]:
        assert bits(port.hgp_sinf(x)) == bits(ref.hgref_sinf(x)), x
        assert bits(port.hgp_cosf(x)) == bits(ref.hgref_cosf(x)), x


def test_transformation_mat4(port, ref):
    rng = np.random.default_rng(2)
    trs = rand_trs(rng, 20000)
    trs[:100, 3:6] *= 100  # large angles -> reduce_large path
    a, b = np.zeros(12, np.float32), np.zeros(12, np.float32)
    for t in trs:
        port.hgp_transformation_mat4(P(t), P(a))
        ref.hgref_transformation_mat4(P(t), P(b))
        assert (bits(a) == bits(b)).all()


def test_mat4_mul_inverse_minmax_visibility(port, ref):
    rng = np.random.default_rng(3)
    fr_p, fr_r = np.zeros(24, np.float32), np.zeros(24, np.float32)
    a, b = np.zeros(12, np.float32), np.zeros(12, np.float32)
    o6a, o6b = np.zeros(6, np.float32), np.zeros(6, np.float32)
    classes = set()
    for it in range(4000):
        m1, m2 = np.zeros(12, np.float32), np.zeros(12, np.float32)
        port.hgp_transformation_mat4(P(rand_trs(rng, 1)[0]), P(m1))
        port.hgp_transformation_mat4(P(rand_trs(rng, 1)[0]), P(m2))
        port.hgp_mat4_mul(P(m1), P(m2), P(a)); ref.hgref_mat4_mul(P(m1), P(m2), P(b))
        assert (bits(a) == bits(b)).all()
        world = a.copy()
        port.hgp_inverse_fast(P(m1), P(a)); ref.hgref_inverse_fast(P(m1), P(b))
        assert (bits(a) == bits(b)).all()
        lo = rng.uniform(-3, 0, 3).astype(np.float32); hi = (lo + rng.uniform(0, 5, 3)).astype(np.float32)
        mm = np.concatenate([lo, hi]).astype(np.float32)
        port.hgp_minmax_transform(P(world), P(mm), P(o6a)); ref.hgref_minmax_transform(P(world), P(mm), P(o6b))
        assert (bits(o6a) == bits(o6b)).all()
        cam = np.zeros(12, np.float32)
        ct = rand_trs(rng, 1)[0]; ct[6:] = 1
        port.hgp_transformation_mat4(P(ct), P(cam))
        args = (np.float32(rng.uniform(0.01, 1)), np.float32(rng.uniform(50, 1000)), np.float32(rng.uniform(0.2, 2.5)),
                np.float32(rng.uniform(0.5, 2)), np.float32(1))
        port.hgp_make_perspective_frustum(*args, P(cam), P(fr_p)); ref.hgref_make_perspective_frustum(*args, P(cam), P(fr_r))
        assert (bits(fr_p) == bits(fr_r)).all()
        va, vb = port.hgp_test_visibility(P(fr_p), P(o6a)), ref.hgref_test_visibility(P(fr_r), P(o6b))
        assert va == vb
        classes.add(va)
    assert classes == {0, 1, 2}


def test_quat_to_euler(port, ref):
    rng = np.random.default_rng(4)
    a, b = np.zeros(3, np.float32), np.zeros(3, np.float32)
    qs = rng.normal(0, 1, (20000, 4)).astype(np.float32)
    qs[0] = 0  # Normalize -> identity
    qs[1] = [0, np.sqrt(.5), 0, np.sqrt(.5)]  # towards the gimbal branch
    qs[2] = [np.sqrt(.5), 0, 0, np.sqrt(.5)]
    qs[3] = [-np.sqrt(.5), 0, 0, np.sqrt(.5)]
    for q in qs:
        port.hgp_quat_to_euler(P(q), P(a)); ref.hgref_quat_to_euler(P(q), P(b))
        assert (bits(a) == bits(b)).all(), q


def test_track_evaluation(port, ref):
    rng = np.random.default_rng(5)
    for it in range(1500):
        n = int(rng.integers(1, 40))
        t = np.sort(rng.choice(np.arange(0, 200), n, replace=False)).astype(np.int64) * 50_000_000
        v = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
        ten = rng.uniform(-.5, .5, n).astype(np.float32); bia = rng.uniform(-.5, .5, n).astype(np.float32)
        q = rng.normal(0, 1, (n, 4)).astype(np.float32)
        for at in [int(t[0]) - 1, int(t[0]), int(t[-1]), int(t[-1]) + 7] + [int(x) for x in rng.integers(-10, 10_000_000_000, 6)] + [int(x) for x in t[:3]]:
            oa, ob = np.full(3, 7, np.float32), np.full(3, 7, np.float32)
            kf = np.zeros(3, np.int32)
            ra = port.hgp_eval_vec3_track(n, P(t), P(v), P(ten), P(bia), at, P(oa))
            rb = ref.hgref_eval_vec3_track(n, P(t), P(v), P(ten), P(bia), at, P(ob), P(kf))
            assert ra == rb and (bits(oa) == bits(ob)).all()
            k0, k1 = C.c_int(-1), C.c_int(-1)
            inside = port.hgp_get_interval_keys(n, P(t), at, C.byref(k0), C.byref(k1))
            assert inside == kf[2] and k0.value == kf[0] and (not inside or k1.value == kf[1])
            qa, qb = np.zeros(4, np.float32), np.zeros(4, np.float32)
            port.hgp_eval_quat_track(n, P(t), P(q), at, P(qa)); ref.hgref_eval_quat_track(n, P(t), P(q), at, P(qb))
            assert (bits(qa) == bits(qb)).all()
    # empty track leaves the value untouched and reports false (animation.h:156-157)
    o = np.full(3, 7, np.float32)
    assert port.hgp_eval_vec3_track(0, None, None, None, None, 5, P(o)) == 0 and (o == 7).all()


def test_world_matrices_and_cull_scene(port, ref):
    """random forest through the reference's Scene API vs the port's flat functions"""
    rng = np.random.default_rng(6)
    n = 3000
    trs = rand_trs(rng, n); trs[:, :3] *= 0.2
    parent = np.full(n, -1, np.int32)
    for i in range(1, n):  # parents may have a LARGER index than the child (exercises the recursion)
        parent[i] = rng.integers(0, n) if rng.random() < 0.15 else rng.integers(0, i)
    # break cycles the way SetTransformParent would refuse them: keep only edges to smaller index or acyclic ones
    for i in range(n):
        k, seen = parent[i], 0
        while k >= 0 and seen <= n:
            if k == i:
                parent[i] = -1
                break
            k = parent[k]; seen += 1
    disabled = (rng.random(n) < 0.05).astype(np.uint8)
    model = rng.integers(0, 3, n).astype(np.int32)
    h = C.c_void_p(ref.hgref_create())
    bounds = []
    for m in range(3):
        lo = rng.uniform(-2, 0, 3); b = np.concatenate([lo, lo + rng.uniform(.1, 3, 3)]).astype(np.float32)
        bounds.append(b)
        ref.hgref_add_model(h, 1, P(b), P(np.zeros(1, np.uint16)), None, None, 0, None)
    ref.hgref_add_nodes(h, n, P(trs), P(parent), P(disabled), P(model), None, None, None)
    ref.hgref_update(h, 0)
    w_ref = np.zeros((n, 12), np.float32)
    assert ref.hgref_get_worlds(h, P(w_ref), 0) == n
    w_port = np.zeros((n, 12), np.float32)
    port.hgp_compute_world_matrices(n, P(trs), P(parent), None, None, P(w_port))
    assert (bits(w_ref) == bits(w_port)).all()

    counts = np.zeros(4, np.int32)
    ref.hgref_get_model_display_lists(h, P(counts))
    assert counts[0] == int((disabled == 0).sum())
    mtx = np.zeros(counts[0], np.uint32); mdl = np.zeros(counts[0], np.uint16); lst = np.zeros(counts[0], np.uint16)
    ref.hgref_read_display_lists(h, 0, P(mtx), P(mdl), P(lst))
    assert (mtx == np.nonzero(disabled == 0)[0]).all()
    cam = np.zeros(12, np.float32); port.hgp_transformation_mat4(P(np.array([0, 0, -30, 0, 0, 0, 1, 1, 1], np.float32)), P(cam))
    fr = np.zeros(24, np.float32); port.hgp_make_perspective_frustum(0.1, 500, 1.0, 16 / 9, 1, P(cam), P(fr))
    pos = np.zeros(counts[0], np.uint32)
    k_ref = ref.hgref_cull(h, 0, P(fr), P(pos), None)
    vis = np.zeros(counts[0], np.uint32)
    k_port = port.hgp_cull(int(counts[0]), P(mtx), P(mdl.astype(np.uint32)), P(np.stack(bounds)), P(w_port), P(fr), P(vis))
    assert k_ref == k_port and 0 < k_ref < counts[0]
    assert (pos[:k_ref] == vis[:k_port]).all()

    # Scene::GetMinMax (scene.cpp:1982) = union over object nodes (disabled ones too) of world * bounds, restated with the port's Arvo
    # transform; min / max are exact, so the boxes agree as values (Union keeps whichever signed zero came first)
    mm_ref = np.zeros(6, np.float32)
    assert ref.hgref_minmax(h, 0, -1, P(mm_ref)) == 1
    boxes = np.zeros((n, 6), np.float32)
    for i in range(n):
        port.hgp_minmax_transform(P(np.ascontiguousarray(w_port[i])), P(bounds[model[i]]), P(boxes[i]))
    assert (np.concatenate([boxes[:, :3].min(0), boxes[:, 3:].max(0)]) == mm_ref).all()
    assert ref.hgref_minmax(h, 100, 50, P(mm_ref)) == 1  # GetNodesMinMax over a node range
    assert (np.concatenate([boxes[100:150, :3].min(0), boxes[100:150, 3:].max(0)]) == mm_ref).all()
    ref.hgref_destroy(h)


def test_reference_camera_view_state(port, ref):
    """The reference's own Scene::ComputeCurrentCameraViewState (scene.cpp:240-263; Compute*ViewState render_pipeline.cpp:67-81, which only
    fetch bgfx::getCaps() and never read it) equals, bit for bit, the frustum the port, the host mirror and the flat probe build from the same
    camera world matrix: the camera-driven frusta of the GPU tests are the reference's."""
    from harfang3d_b200 import hostmirror, scenegen
    from tests.scene_io import RefScene
    fs = scenegen.static_scene(800, 4, 5, n_models=2, seed=321)
    a = RefScene(ref, fs)
    a.update(0)
    for node, fov in ((777, 0.9), (5, 2.4), (300, 0.001)):
        assert a.set_camera(node, 0.2, 350.0, fov=fov)
        fr, proj, view = a.view_state(16, 9)
        w = np.ascontiguousarray(a.worlds()[ref.hgref_node_transform_idx(a.h, node)])
        probe, inv = np.zeros(24, np.float32), np.zeros(12, np.float32)
        ref.hgref_make_perspective_frustum(0.2, 350.0, fov, 16, 9, P(w), P(probe))
        ref.hgref_inverse_fast(P(w), P(inv))
        assert (bits(probe) == bits(fr.reshape(-1))).all() and (bits(inv) == bits(view.reshape(-1))).all()
        fp = np.zeros(24, np.float32)
        port.hgp_make_perspective_frustum(0.2, 350.0, fov, 16, 9, P(w), P(fp))
        assert (bits(fp) == bits(probe)).all()
        assert (bits(hostmirror.make_perspective_frustum(0.2, 350.0, fov, 16, 9, w).reshape(-1)) == bits(probe)).all()
        assert proj[3, 2] == 1 and proj[3, 3] == 0 and proj[0, 0] * 16 == pytest.approx(proj[1, 1] * 9, rel=1e-6)
    assert a.set_camera(9, 0.5, 90.0, size=12.0)
    fr, proj, _ = a.view_state(4, 3)
    w = np.ascontiguousarray(a.worlds()[ref.hgref_node_transform_idx(a.h, 9)])
    probe = np.zeros(24, np.float32)
    ref.hgref_make_ortho_frustum(0.5, 90.0, 12.0, 4, 3, P(w), P(probe))
    assert (bits(probe) == bits(fr.reshape(-1))).all() and proj[3, 3] == 1
    a.close()
