"""Generates the golden fixtures of tests/golden/ from the UNMODIFIED reference (oracle/_ref/libhg_ref.so, built from /root/reference by
`make -C oracle ref`). Run in the build container only:   python tests/golden/gen_golden.py
The fixtures travel with the repo so that the oracle port and the GPU path can be checked against the reference's own outputs on
machines where /root/reference does not exist.

  math.npz    inputs + reference outputs of every arithmetic function on the path (bit patterns)
  scenes.npz  reference frames of small seeded scenes (scenegen): worlds, TRS after animation, visible lists, bone matrices
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from harfang3d_b200 import hostmirror, scenegen  # noqa: E402
from tests import oracle_lib  # noqa: E402
from tests.oracle_lib import P  # noqa: E402
from tests.scene_io import RefScene  # noqa: E402


def golden_scenes():
    """name -> (FlatScene factory, frames, dt). Shared with the tests so both sides build the same scenes."""
    def pos_tracks():
        rng = np.random.default_rng(41)
        fs = scenegen.static_scene(5_000, 6, 6, n_models=4, seed=41)
        scenegen.add_position_tracks(fs, np.sort(rng.choice(fs.n, 1_500, replace=False)).astype(np.int32), rng, n_keys=16)
        return fs
    return {
        "c1": (lambda: scenegen.config_c1(), 1, 0),
        "c3s": (lambda: scenegen.config_c3(n_chars=8, n_bones=64, n_keys=32), 5, 50_000_000),
        "pos": (pos_tracks, 4, scenegen.FRAME_DT),
        "c5s": (lambda: scenegen.config_c5(n_instances=4), 1, 0),
    }


def scene_digest(fs):
    h = hashlib.sha256()
    for a in (fs.trs, fs.parent, fs.disabled, fs.model, fs.v3_val, fs.q_val, fs.v3_t):
        h.update(np.ascontiguousarray(a).tobytes())
    return np.frombuffer(h.digest()[:8], np.uint64)[0]


def gen_math(ref):
    rng = np.random.default_rng(2024)
    n = 1500
    out = {}
    trs = np.empty((n, 9), np.float32)
    trs[:, 0:3] = rng.uniform(-50, 50, (n, 3)); trs[:, 3:6] = rng.uniform(-np.pi, np.pi, (n, 3)); trs[:, 6:9] = rng.uniform(0.5, 2, (n, 3))
    trs[:40, 3:6] *= 300  # |angle| >= 120: large-argument reduction of sinf/cosf
    trs[40:60, 3:6] *= 1e-5
    m = np.zeros((n, 12), np.float32)
    for i in range(n):
        ref.hgref_transformation_mat4(P(trs[i]), P(m[i]))
    out["trs"], out["trs_mat4"] = trs, m
    mul, inv = np.zeros((n, 12), np.float32), np.zeros((n, 12), np.float32)
    for i in range(n):
        ref.hgref_mat4_mul(P(m[i]), P(m[(i * 7 + 1) % n]), P(mul[i]))
        ref.hgref_inverse_fast(P(m[i]), P(inv[i]))
    out["mat4_mul"], out["inverse_fast"] = mul, inv
    lo = rng.uniform(-3, 0, (n, 3)); mm = np.concatenate([lo, lo + rng.uniform(0, 5, (n, 3))], axis=1).astype(np.float32)
    mm_out = np.zeros((n, 6), np.float32)
    for i in range(n):
        ref.hgref_minmax_transform(P(mul[i]), P(mm[i]), P(mm_out[i]))
    out["minmax_in"], out["minmax_out"] = mm, mm_out
    cam_trs = trs.copy(); cam_trs[:, 6:9] = 1
    fr_args = np.stack([rng.uniform(0.01, 1, n), rng.uniform(50, 1000, n), rng.uniform(0.2, 2.5, n), rng.uniform(0.5, 2, n), np.ones(n)], axis=1).astype(np.float32)
    fr, vis = np.zeros((n, 24), np.float32), np.zeros(n, np.int32)
    for i in range(n):
        cam = np.zeros(12, np.float32)
        ref.hgref_transformation_mat4(P(cam_trs[(i * 3) % n]), P(cam))
        ref.hgref_make_perspective_frustum(*[np.float32(x) for x in fr_args[i]], P(cam), P(fr[i]))
        vis[i] = ref.hgref_test_visibility(P(fr[i]), P(mm_out[i]))
    out["frustum_cam_trs"], out["frustum_args"], out["frustum"], out["visibility"] = cam_trs, fr_args, fr, vis
    q = rng.normal(0, 1, (n, 4)).astype(np.float32)
    q[0] = 0; q[1] = [0, np.sqrt(.5), 0, np.sqrt(.5)]; q[2] = [np.sqrt(.5), 0, 0, np.sqrt(.5)]; q[3] = [-np.sqrt(.5), 0, 0, np.sqrt(.5)]
    e = np.zeros((n, 3), np.float32)
    for i in range(n):
        ref.hgref_quat_to_euler(P(q[i]), P(e[i]))
    out["quat"], out["quat_euler"] = q, e
    xs = np.concatenate([rng.uniform(-np.pi, np.pi, 4000), rng.uniform(-500, 500, 1000), rng.normal(0, 1e-3, 500), [0.0, -0.0, 1e30, -1e30, 119.99, 120.01]]).astype(np.float32)
    out["sincos_x"] = xs
    out["sinf"] = np.array([ref.hgref_sinf(float(x)) for x in xs], np.float32)
    out["cosf"] = np.array([ref.hgref_cosf(float(x)) for x in xs], np.float32)
    # tracks: 200 random tracks x 12 times
    K = 12
    tk_n, tk_t, tk_v, tk_ten, tk_bias, tk_q, tk_at, tk_out3, tk_ret, tk_kf, tk_outq = [], [], [], [], [], [], [], [], [], [], []
    for it in range(200):
        nk = int(rng.integers(1, 24))
        t = np.zeros(24, np.int64); t[:nk] = np.sort(rng.choice(np.arange(0, 200), nk, replace=False)).astype(np.int64) * 50_000_000
        v = np.zeros((24, 3), np.float32); v[:nk] = rng.uniform(-10, 10, (nk, 3))
        ten = np.zeros(24, np.float32); ten[:nk] = rng.uniform(-.5, .5, nk)
        bia = np.zeros(24, np.float32); bia[:nk] = rng.uniform(-.5, .5, nk)
        qq = np.zeros((24, 4), np.float32); qq[:nk] = rng.normal(0, 1, (nk, 4))
        ats = np.array([t[0] - 1, t[0], t[nk - 1], t[nk - 1] + 7] + list(rng.integers(-10, 10_000_000_000, K - 6)) + [t[min(1, nk - 1)], t[nk // 2]], np.int64)
        o3, oq, rets, kfs = np.zeros((K, 3), np.float32), np.zeros((K, 4), np.float32), np.zeros(K, np.int32), np.zeros((K, 3), np.int32)
        for j, at in enumerate(ats):
            o3[j] = 7
            rets[j] = ref.hgref_eval_vec3_track(nk, P(t), P(v), P(ten), P(bia), int(at), P(o3[j]), P(kfs[j]))
            ref.hgref_eval_quat_track(nk, P(t), P(qq), int(at), P(oq[j]))
        tk_n.append(nk); tk_t.append(t); tk_v.append(v); tk_ten.append(ten); tk_bias.append(bia); tk_q.append(qq); tk_at.append(ats)
        tk_out3.append(o3); tk_ret.append(rets); tk_kf.append(kfs); tk_outq.append(oq)
    for k, v in dict(tk_n=tk_n, tk_t=tk_t, tk_v=tk_v, tk_ten=tk_ten, tk_bias=tk_bias, tk_q=tk_q, tk_at=tk_at, tk_out3=tk_out3, tk_ret=tk_ret, tk_kf=tk_kf,
                     tk_outq=tk_outq).items():
        out[k] = np.array(v)
    return out


def gen_scenes(ref):
    out = {}
    for name, (make, frames, dt) in golden_scenes().items():
        fs = make()
        out[name + "_digest"] = scene_digest(fs)
        rs = RefScene(ref, fs)
        for _ in range(frames):
            rs.update(dt)
        out[name + "_worlds"] = rs.worlds()[: fs.n]
        out[name + "_worlds_prev"] = rs.worlds(previous=True)[: fs.n]
        counts, lists = rs.display_lists()
        out[name + "_dl_counts"] = counts
        out[name + "_dl_mtx"] = lists[0][0]
        if len(fs.anim_node):
            nodes = np.unique(fs.anim_node)
            out[name + "_trs_nodes"] = nodes
            out[name + "_trs"] = rs.trs(nodes)
        frusta = hostmirror.scene_frusta(fs)
        for v in range(len(frusta)):
            pos, mtx = rs.cull(0, frusta[v])
            out["%s_vis%d_pos" % (name, v)] = pos
            out["%s_vis%d_mtx" % (name, v)] = mtx
        if fs.bones:
            out[name + "_bones"] = rs.bone_matrices()
            out[name + "_bones_prev"] = rs.bone_matrices(previous=True)
        rs.close()
    return out


if __name__ == "__main__":
    ref = oracle_lib.load_ref()
    assert ref is not None, "build oracle/_ref first: make -C oracle ref"
    np.savez_compressed(os.path.join(HERE, "math.npz"), **gen_math(ref))
    np.savez_compressed(os.path.join(HERE, "scenes.npz"), **gen_scenes(ref))
    for f in ("math.npz", "scenes.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
