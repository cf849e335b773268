"""TEST INFRASTRUCTURE: ctypes loaders for the two oracles.

  load_ref()  -> oracle/_ref/libhg_ref.so : the unmodified reference compiled from /root/reference (may be absent)
  load_port() -> oracle/libhg_port.so     : plain-C restatement, built on demand with gcc

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE = os.path.join(ROOT, "oracle")

c_f, c_i, c_i64, c_p = C.c_float, C.c_int, C.c_int64, C.c_void_p


def P(a):
    """pointer to a C-contiguous numpy array (or None)"""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "array must be contiguous"
    return a.ctypes.data_as(C.c_void_p)


def _sig(lib, name, restype, argtypes):
    f = getattr(lib, name)
    f.restype, f.argtypes = restype, argtypes


def load_port():
    so = os.path.join(ORACLE, "libhg_port.so")
    src = [os.path.join(ORACLE, f) for f in ("hg_port.c", "hg_port.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", ORACLE, "port"], stdout=subprocess.DEVNULL)
    L = C.CDLL(so)
    _sig(L, "hgp_sinf", c_f, [c_f])
    _sig(L, "hgp_cosf", c_f, [c_f])
    for n in ("hgp_transformation_mat4", "hgp_inverse_fast", "hgp_quat_to_euler", "hgp_to_bgfx"):
        _sig(L, n, None, [c_p, c_p])
    for n in ("hgp_mat4_mul", "hgp_minmax_transform"):
        _sig(L, n, None, [c_p, c_p, c_p])
    _sig(L, "hgp_test_visibility", c_i, [c_p, c_p])
    _sig(L, "hgp_make_perspective_frustum", None, [c_f] * 5 + [c_p, c_p])
    _sig(L, "hgp_get_interval_keys", c_i, [c_i, c_p, c_i64, c_p, c_p])
    _sig(L, "hgp_eval_vec3_track", c_i, [c_i, c_p, c_p, c_p, c_p, c_i64, c_p])
    _sig(L, "hgp_eval_quat_track", c_i, [c_i, c_p, c_p, c_i64, c_p])
    _sig(L, "hgp_step_play_clock", c_i, [c_p, c_i64, C.c_int8, c_i, c_i64, c_i64, c_i])
    _sig(L, "hgp_compute_world_matrices", None, [c_i, c_p, c_p, c_p, c_p, c_p])
    _sig(L, "hgp_cull", c_i, [c_i, c_p, c_p, c_p, c_p, c_p, c_p])
    _sig(L, "hgp_bone_matrices", None, [c_i, c_p, c_p, c_p, c_p, c_p])
    return L


def _bind_scene_driver(L, with_probes):
    from harfang3d_b200 import flatscene
    return flatscene.bind(L, with_probes)


def load_ref():
    so = os.path.join(ORACLE, "_ref", "libhg_ref.so")
    if not os.path.exists(so):
        return None
    return _bind_scene_driver(C.CDLL(so), True)


def load_facade_driver():
    """the flat binding compiled against the PRODUCT's hg::Scene facade (harfang3d_b200/libhg_facade_flat.so): same flat calls, so
    tests.scene_io.RefScene drives the facade exactly like it drives the reference. None when not built."""
    from harfang3d_b200 import flatscene
    try:
        return flatscene.load()
    except RuntimeError:
        return None


def bits(a):
    """bit pattern view for exact float comparison (distinguishes -0.0, compares NaN payloads)"""
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
