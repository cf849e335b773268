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
    """signatures of oracle/ref_driver.cpp (same file for the reference build and the facade build)"""
    _sig(L, "hgref_create", c_p, [])
    _sig(L, "hgref_destroy", None, [c_p])
    _sig(L, "hgref_add_model", c_i, [c_p, c_i, c_p, c_p, c_p, c_p, c_i, c_p])
    _sig(L, "hgref_add_nodes", c_i, [c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_set_object_bones", c_i, [c_p, c_i, c_i, c_p])
    _sig(L, "hgref_set_trs", None, [c_p, c_i, c_p])
    _sig(L, "hgref_get_trs", None, [c_p, c_i, c_p])
    _sig(L, "hgref_set_parent", None, [c_p, c_i, c_i])
    _sig(L, "hgref_enable_node", None, [c_p, c_i, c_i])
    _sig(L, "hgref_destroy_node", None, [c_p, c_i])
    _sig(L, "hgref_garbage_collect", c_i, [c_p])
    _sig(L, "hgref_set_node_world", None, [c_p, c_i, c_p])
    _sig(L, "hgref_set_transform_matrix", None, [c_p, c_i, c_p, c_i])
    _sig(L, "hgref_minmax", c_i, [c_p, c_i, c_i, c_p])
    _sig(L, "hgref_load_binary", c_i, [c_p, C.c_char_p])
    if with_probes:
        _sig(L, "hgref_save_binary", c_i, [c_p, C.c_char_p])
    _sig(L, "hgref_node_transform_idx", c_i, [c_p, c_i])
    _sig(L, "hgref_add_anim", c_i, [c_p, c_i])
    _sig(L, "hgref_anim_add_vec3_track", c_i, [c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_anim_add_quat_track", c_i, [c_p, c_i, c_i, c_p, c_p])
    _sig(L, "hgref_anim_add_enable_track", c_i, [c_p, c_i, c_i, c_p, c_p])
    _sig(L, "hgref_anim_add_float_track", c_i, [c_p, c_i, C.c_char_p, c_i, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_is_node_enabled", c_i, [c_p, c_i])
    _sig(L, "hgref_add_anims_bulk", c_i, [c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_add_scene_anim", c_i, [c_p, c_i, c_p, c_p, c_i64, c_i64])
    _sig(L, "hgref_play_anim", c_i, [c_p, c_i, c_i, c_f])
    _sig(L, "hgref_is_playing", c_i, [c_p, c_i])
    _sig(L, "hgref_update", None, [c_p, c_i64])
    _sig(L, "hgref_update_phase", None, [c_p, c_i, c_i64])
    _sig(L, "hgref_world_count", c_i, [c_p])
    _sig(L, "hgref_get_worlds", c_i, [c_p, c_p, c_i])
    _sig(L, "hgref_get_model_display_lists", None, [c_p, c_p])
    _sig(L, "hgref_read_display_lists", c_i, [c_p, c_i, c_p, c_p, c_p])
    _sig(L, "hgref_read_skinned_lists", c_i, [c_p, c_i, c_p, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_cull", c_i, [c_p, c_i, c_p, c_p, c_p])
    _sig(L, "hgref_bone_matrices", c_i, [c_p, c_i, c_i, c_p])
    _sig(L, "hgref_set_node_camera", c_i, [c_p, c_i, c_i, c_f, c_f, c_f])
    _sig(L, "hgref_set_camera_fov", None, [c_p, c_i, c_f])
    _sig(L, "hgref_get_camera_fov", c_f, [c_p, c_i])
    _sig(L, "hgref_set_current_camera", None, [c_p, c_i])
    _sig(L, "hgref_current_camera_view_state", None, [c_p, c_f, c_f, c_p])
    _sig(L, "hgref_make_perspective_frustum", None, [c_f] * 5 + [c_p, c_p])
    _sig(L, "hgref_make_ortho_frustum", None, [c_f] * 5 + [c_p, c_p])
    for n in ("hgref_transformation_mat4", "hgref_inverse_fast"):
        _sig(L, n, None, [c_p, c_p])
    _sig(L, "hgref_mat4_mul", None, [c_p, c_p, c_p])
    _sig(L, "hgref_sinf", c_f, [c_f])
    _sig(L, "hgref_cosf", c_f, [c_f])
    _sig(L, "hgref_time_frames", c_i, [c_p, c_i, c_i64, c_p, c_i, c_p])
    if with_probes:  # reference-internal host math: compiled out of the facade build
        _sig(L, "hgref_minmax_transform", None, [c_p, c_p, c_p])
        _sig(L, "hgref_quat_to_euler", None, [c_p, c_p])
        _sig(L, "hgref_test_visibility", c_i, [c_p, c_p])
        _sig(L, "hgref_sort_keys", c_i, [c_p, c_p, c_p])
        _sig(L, "hgref_eval_vec3_track", c_i, [c_i, c_p, c_p, c_p, c_p, c_i64, c_p, c_p])
        _sig(L, "hgref_eval_quat_track", c_i, [c_i, c_p, c_p, c_i64, c_p])
    return L


def load_ref():
    so = os.path.join(ORACLE, "_ref", "libhg_ref.so")
    if not os.path.exists(so):
        return None
    return _bind_scene_driver(C.CDLL(so), True)


def load_facade_driver():
    """oracle/ref_driver.cpp compiled against the PRODUCT's hg::Scene facade (make -C oracle facade_driver): same flat calls, so
    tests.scene_io.RefScene drives the facade exactly like it drives the reference. None when not built."""
    so = os.path.join(ORACLE, "_facade", "libhg_facade_driver.so")
    if not os.path.exists(so):
        return None
    return _bind_scene_driver(C.CDLL(so), False)


def bits(a):
    """bit pattern view for exact float comparison (distinguishes -0.0, compares NaN payloads)"""
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
