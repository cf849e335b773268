"""ctypes table of the flat C binding over the hg::Scene API (csrc/facade/flat/hg_flat.cpp -> libhg_facade_flat.so): build a scene from
a scenegen.FlatScene, run frames, read worlds / display lists / visible lists / bone matrices back. bench.py's `e2e_facade` leg and the
facade tests drive the product's hg::Scene facade through it; the test suite binds the same entry points of its reference build with
`bind()` (the binding only uses the public hg::Scene interface, so one table serves both).
"""
import ctypes as C
import os

import numpy as np

c_f, c_i, c_i64, c_p = C.c_float, C.c_int, C.c_int64, C.c_void_p
HERE = os.path.dirname(os.path.abspath(__file__))


def P(a):
    """pointer to a C-contiguous numpy array (or None)"""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "array must be contiguous"
    return a.ctypes.data_as(C.c_void_p)


def _sig(lib, name, restype, argtypes):
    f = getattr(lib, name)
    f.restype, f.argtypes = restype, argtypes


def bind(L, with_probes):
    """signatures of csrc/facade/flat/hg_flat.cpp (the same file binds any implementation of the hg::Scene API)"""
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
    _sig(L, "hgref_add_point_light", c_i, [c_p, c_i, c_f])
    _sig(L, "hgref_get_light", None, [c_p, c_i, c_p])
    _sig(L, "hgref_anim_add_color_track", c_i, [c_p, c_i, C.c_char_p, c_i, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_anim_add_vec4_track", c_i, [c_p, c_i, C.c_char_p, c_i, c_p, c_p, c_p, c_p])
    _sig(L, "hgref_set_material_value", None, [c_p, c_i, c_i, C.c_char_p, c_p])
    _sig(L, "hgref_get_material_value", c_i, [c_p, c_i, c_i, C.c_char_p, c_p])
    _sig(L, "hgref_set_scene_anim_name", None, [c_p, c_i, C.c_char_p])
    _sig(L, "hgref_add_instance", c_i, [c_p, c_i, C.c_char_p, C.c_char_p, c_i])
    _sig(L, "hgref_setup_instance", c_i, [c_p, c_i])
    _sig(L, "hgref_node_count", c_i, [c_p])
    _sig(L, "hgref_anim_set_instance_track", c_i, [c_p, c_i, c_i, c_p, C.c_char_p, c_p, c_p])
    _sig(L, "hgref_time_frames", c_i, [c_p, c_i, c_i64, c_p, c_i, c_p])
    if with_probes:  # reference-internal host math: compiled out of the facade build
        _sig(L, "hgref_minmax_transform", None, [c_p, c_p, c_p])
        _sig(L, "hgref_quat_to_euler", None, [c_p, c_p])
        _sig(L, "hgref_test_visibility", c_i, [c_p, c_p])
        _sig(L, "hgref_sort_keys", c_i, [c_p, c_p, c_p])
        _sig(L, "hgref_eval_vec3_track", c_i, [c_i, c_p, c_p, c_p, c_p, c_i64, c_p, c_p])
        _sig(L, "hgref_eval_quat_track", c_i, [c_i, c_p, c_p, c_i64, c_p])
    return L



def load():
    """the product's build: the flat binding over the hg::Scene facade (make -C harfang3d_b200 facade_flat). Raises when it is missing."""
    so = os.path.join(HERE, "libhg_facade_flat.so")
    if not os.path.exists(so):
        raise RuntimeError("libhg_facade_flat.so is not built (python -c 'import __graft_entry__ as g; g.build()')")
    L = bind(C.CDLL(so), False)
    _sig(L, "hgfacade_time_fused_frames", c_i, [c_p, c_i, c_i64, c_p, c_p])
    _sig(L, "hgfacade_read_visible", c_i, [c_p, c_p, c_i])
    _sig(L, "hgfacade_set_default_device", None, [c_i])
    _sig(L, "hgfacade_visible_breakdown", None, [c_p])
    _sig(L, "hgfacade_set_default_append_slack", None, [c_i])
    _sig(L, "hgfacade_launch_count", C.c_uint64, [c_p])
    _sig(L, "hgfacade_rebuild_count", C.c_uint64, [c_p])
    return L


class SceneHandle:
    """a scene built from a FlatScene through the flat binding, plus the per-frame calls"""
    def __init__(self, ref, fs, binary=None):
        self.L, self.fs = ref, fs
        self.h = C.c_void_p(ref.hgref_create())
        for m in fs.models:
            L_ = m["bounds"].shape[0]
            if m["bone_tables"] is not None:
                off = np.cumsum([0] + [len(t) for t in m["bone_tables"]]).astype(np.int32)
                tbl = np.concatenate(m["bone_tables"]).astype(np.uint16)
            else:
                off, tbl = None, None
            bp = m["bind_pose"]
            ref.hgref_add_model(self.h, L_, P(np.ascontiguousarray(m["bounds"], np.float32)), P(np.ascontiguousarray(m["mats"], np.uint16)),
                                P(off), P(tbl), 0 if bp is None else bp.shape[0], None if bp is None else P(np.ascontiguousarray(bp, np.float32)))
        if binary is not None:
            # the scene comes from a scene-binary file (Scene::Load_binary); models were registered above under the names the file refers to
            n = ref.hgref_load_binary(self.h, os.fsencode(binary))
            assert n == fs.n, "scene binary: %d nodes loaded, %d expected" % (n, fs.n)
        else:
            ref.hgref_add_nodes(self.h, fs.n, P(np.ascontiguousarray(fs.trs)), P(fs.parent), P(fs.disabled), P(fs.model), P(fs.n_mats),
                                P(fs.transparent_mask), None)
            for node, b in fs.bones.items():
                ref.hgref_set_object_bones(self.h, int(node), len(b), P(np.ascontiguousarray(b, np.int32)))
            if len(fs.anim_node):
                c = np.ascontiguousarray
                ref.hgref_add_anims_bulk(self.h, len(fs.anim_node), P(c(fs.anim_use_quat, np.uint8)), len(fs.trk_anim), P(c(fs.trk_anim, np.int32)),
                                         P(c(fs.trk_kind, np.uint8)), P(c(fs.trk_key_off, np.uint32)), P(c(fs.trk_key_count, np.uint32)), P(c(fs.v3_t, np.int64)),
                                         P(c(fs.v3_val, np.float32)), P(c(fs.v3_tension, np.float32)), P(c(fs.v3_bias, np.float32)), P(c(fs.q_t, np.int64)),
                                         P(c(fs.q_val, np.float32)))
            for anims, ts, te in fs.scene_anims:
                nodes = np.ascontiguousarray(fs.anim_node[anims], np.int32)
                ref.hgref_add_scene_anim(self.h, len(anims), P(nodes), P(np.ascontiguousarray(anims, np.int32)), int(ts), int(te))
        for sa, loop, scale in fs.plays:
            ref.hgref_play_anim(self.h, sa, loop, scale)

    def save_binary(self, path):
        """Scene::Save_binary (reference build only)"""
        assert self.L.hgref_save_binary(self.h, os.fsencode(path)) == 1

    def close(self):
        if self.h:
            self.L.hgref_destroy(self.h)
            self.h = None

    def update(self, dt):
        self.L.hgref_update(self.h, int(dt))

    def worlds(self, previous=False):
        n = self.L.hgref_world_count(self.h) if not previous else self.L.hgref_get_worlds(self.h, None, 1)
        w = np.zeros((n, 12), np.float32)
        self.L.hgref_get_worlds(self.h, P(w), 1 if previous else 0)
        return w

    def trs(self, nodes):
        out = np.zeros((len(nodes), 9), np.float32)
        for i, nd in enumerate(nodes):
            self.L.hgref_get_trs(self.h, int(nd), P(out[i]))
        return out

    def display_lists(self):
        """returns counts and (mtx_idx, mdl_idx, lst_idx) of the opaque and transparent lists"""
        counts = np.zeros(4, np.int32)
        self.L.hgref_get_model_display_lists(self.h, P(counts))
        out = {}
        for kind in (0, 1):
            n = int(counts[kind])
            mtx, mdl, lst = np.zeros(n, np.uint32), np.zeros(n, np.uint16), np.zeros(n, np.uint16)
            self.L.hgref_read_display_lists(self.h, kind, P(mtx), P(mdl), P(lst))
            out[kind] = (mtx, mdl, lst)
        return counts, out

    def cull(self, kind, frustum24):
        """copy + CullModelDisplayLists; returns the surviving (mtx_idx) sequence and positions in the un-culled list"""
        counts = np.zeros(4, np.int32)
        self.L.hgref_get_model_display_lists(self.h, P(counts))
        n = int(counts[kind])
        pos, mtx = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.uint32)
        k = self.L.hgref_cull(self.h, kind, P(np.ascontiguousarray(frustum24, np.float32)), P(pos), P(mtx))
        return pos[:k].copy(), mtx[:k].copy()

    def skinned_lists(self, kind=2):
        counts = np.zeros(4, np.int32)
        self.L.hgref_get_model_display_lists(self.h, P(counts))
        n = int(counts[kind])
        mi, bi = np.zeros((max(n, 1), 32), np.uint32), np.zeros((max(n, 1), 32), np.uint32)
        bc, mdl, lst = np.zeros(max(n, 1), np.uint16), np.zeros(max(n, 1), np.uint16), np.zeros(max(n, 1), np.uint16)
        self.L.hgref_read_skinned_lists(self.h, kind, P(mi), P(bi), P(bc), P(mdl), P(lst))
        return mi[:n], bi[:n], bc[:n], mdl[:n], lst[:n]

    def set_transform_matrix(self, node, m12, local=False):
        """Scene::SetTransformWorldMatrix / SetTransformLocalMatrix"""
        self.L.hgref_set_transform_matrix(self.h, int(node), P(np.ascontiguousarray(m12, np.float32).reshape(12)), 1 if local else 0)

    def set_camera(self, node, znear, zfar, fov=None, size=None, current=True):
        """Scene::CreateCamera / CreateOrthographicCamera + SetNodeCamera (+ SetCurrentCamera)"""
        ok = self.L.hgref_set_node_camera(self.h, int(node), 1 if size is not None else 0, float(znear), float(zfar), float(size if size is not None else fov))
        if current:
            self.L.hgref_set_current_camera(self.h, int(node))
        return bool(ok)

    def view_state(self, ar_x, ar_y):
        """Scene::ComputeCurrentCameraViewState -> (frustum[6,4], proj[4,4], view[3,4])"""
        out = np.zeros(52, np.float32)
        self.L.hgref_current_camera_view_state(self.h, float(ar_x), float(ar_y), P(out))
        return out[:24].reshape(6, 4).copy(), out[24:40].reshape(4, 4).copy(), out[40:].reshape(3, 4).copy()

    def minmax(self, first=0, count=-1):
        """Scene::GetMinMax (count < 0) or GetNodesMinMax over nodes [first, first+count) -> (ok, minmax6)"""
        out = np.zeros(6, np.float32)
        ok = self.L.hgref_minmax(self.h, int(first), int(count), P(out))
        return bool(ok), out

    def sort_keys(self, view12):
        """ComputeModelDisplayListSortKeys over the list left by the last cull()"""
        v = np.ascontiguousarray(view12, np.float32).reshape(12)
        n = self.L.hgref_sort_keys(self.h, P(v), None)
        out = np.zeros(max(n, 1), np.uint32)
        self.L.hgref_sort_keys(self.h, P(v), P(out))
        return out[:n]

    def bone_matrices(self, previous=False, kind=2):
        counts = np.zeros(4, np.int32)
        self.L.hgref_get_model_display_lists(self.h, P(counts))
        n = self.L.hgref_bone_matrices(self.h, kind, 1 if previous else 0, None)
        out = np.zeros((n, 16), np.float32)
        self.L.hgref_bone_matrices(self.h, kind, 1 if previous else 0, P(out))
        return out
