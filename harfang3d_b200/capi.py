"""ctypes binding of include/hgb200.h. One method per C entry point, same names and argument meaning; errors raise HgbError
with hgb_last_error(). Arrays are numpy, C-contiguous, exact dtype (checked) — nothing is converted silently."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = 0xFFFFFFFF
NOT_PLAYING = -(2**63)
ENTRY_DISABLED = 0x01
KIND_OPAQUE, KIND_TRANSPARENT = 0, 1


KERNEL_CLASSES = ("k_sample_tracks", "k_update_level", "k_compact", "k_segment_ranks", "k_cull_flags", "k_bone_matrices", "k_sort_keys", "k_minmax")


class HgbError(RuntimeError):
    pass


def library_path():
    return os.path.join(_HERE, "libhgb200.so")


class Track(C.Structure):
    _fields_ = [("target_slot", C.c_uint32), ("key_off", C.c_uint32), ("key_count", C.c_uint32), ("play", C.c_uint16), ("kind", C.c_uint8),
                ("pad", C.c_uint8)]


TRACK_DTYPE = np.dtype([("target_slot", "<u4"), ("key_off", "<u4"), ("key_count", "<u4"), ("play", "<u2"), ("kind", "u1"), ("pad", "u1")])

_u32, _i, _p, _sz = C.c_uint32, C.c_int, C.c_void_p, C.c_size_t

# name -> (restype, argtypes). Kept in one table so tests can check that every symbol of include/hgb200.h is exported.
SIGNATURES = {
    "hgb_last_error": (C.c_char_p, []),
    "hgb_device_count": (_i, []),
    "hgb_create": (_p, [_i, _p]),
    "hgb_destroy": (None, [_p]),
    "hgb_set_stream": (_i, [_p, _p]),
    "hgb_sync": (_i, [_p]),
    "hgb_set_topology": (_i, [_p, _u32, _p, _p]),
    "hgb_upload_trs": (_i, [_p, _u32, _u32, _p]),
    "hgb_set_parents": (_i, [_p, _u32, _p, _p, _p]),
    "hgb_compact_visible_noalloc": (_i, [_i, _p, _u32, _u32, _p, _u32, _p, _p]),
    "hgb_append_nodes": (_i, [_p, _u32, _p, _p, _p, _p]),
    "hgb_slot_capacity": (_i, [_p, _p]),
    "hgb_remove_nodes": (_i, [_p, _u32, _p, _p]),
    "hgb_append_display_entries": (_i, [_p, _i, _u32, _p, _p, _p, _p, _p]),
    "hgb_set_preset_world": (_i, [_p, _u32, _p, _p]),
    "hgb_set_matrices": (_i, [_p, _u32, _p, _p, _i]),
    "hgb_set_bounds": (_i, [_p, _u32, _p]),
    "hgb_set_display_entries": (_i, [_p, _i, _u32, _p, _p, _p, _p]),
    "hgb_set_entry_flags": (_i, [_p, _i, _u32, _u32, _p]),
    "hgb_set_skins": (_i, [_p, _u32, _p, _p, _u32, _p]),
    "hgb_set_anim_keys": (_i, [_p, _u32, _p, _p, _p, _p, _u32, _p, _p]),
    "hgb_set_tracks": (_i, [_p, _u32, _p, _u32]),
    "hgb_set_views": (_i, [_p, _u32, _p]),
    "hgb_set_instance_views": (_i, [_p, _u32, _p, _p]),
    "hgb_update": (_i, [_p, _u32, _p, _i]),
    "hgb_cull": (_i, [_p]),
    "hgb_cull_list": (_i, [_p, _u32, _p, _p, _u32, _p, _p, _p, _p]),
    "hgb_bone_matrices": (_i, [_p, _i]),
    "hgb_sort_keys": (_i, [_p, _i, _u32, _p]),
    "hgb_cull_skinned_list": (_i, [_p, _u32, _p, _p, _p, _u32, _p, _p, _u32, _p, _p, _p, _p]),
    "hgb_set_entry_records": (_i, [_p, _i, _u32, _u32, _p]),
    "hgb_get_visible_records_ptr": (_i, [_p, _i, _u32, _p, _p]),
    "hgb_get_visible_records_begin": (_i, [_p, _i, _u32, _u32, _p, _p]),
    "hgb_visible_records_wait": (_i, [_p, _i, _u32]),
    "hgb_get_visible_ptr": (_i, [_p, _i, _u32, _p, _p]),
    "hgb_peer_init": (_i, [_p, _u32, _u32, _u32, _p, _p, _p, _p]),
    "hgb_peer_connect": (_i, [_p, _p, _p]),
    "hgb_peer_close": (_i, [_p]),
    "hgb_peer_status": (_i, [_p, _p, _p]),
    "hgb_peer_result": (_i, [_p, _p, _p, _u32, _p, _u32]),
    "hgb_set_enable_tracks": (_i, [_p, _u32, _p, _u32, _p, _p]),
    "hgb_set_enable_state": (_i, [_p, _u32, _p, _p]),
    "hgb_get_enable_state": (_i, [_p, _u32, _p, _p, _p]),
    "hgb_set_objects": (_i, [_p, _u32, _p, _p, _u32, _p, _p, _u32]),
    "hgb_minmax": (_i, [_p]),
    "hgb_get_minmax": (_i, [_p, _p, _p]),
    "hgb_get_sort_keys": (_i, [_p, _i, _u32, _p, _u32, _p]),
    "hgb_get_worlds": (_i, [_p, _u32, _u32, _p, _i]),
    "hgb_get_trs": (_i, [_p, _u32, _u32, _p]),
    "hgb_get_visible_count": (_i, [_p, _i, _u32, _p]),
    "hgb_get_visible": (_i, [_p, _i, _u32, _p, _u32, _p]),
    "hgb_get_instance_offsets": (_i, [_p, _p]),
    "hgb_get_bone_matrices": (_i, [_p, _p, _i]),
    "hgb_device_ptr": (_p, [_p, _i]),
    "hgb_launch_count": (C.c_uint64, [_p]),
    "hgb_rebuild_count": (C.c_uint64, [_p]),
    "hgb_frame_was_async": (_i, [_p]),
    "hgb_level_count": (_u32, [_p]),
    "hgb_is_identity_order": (_i, [_p]),
    "hgb_time_begin": (_i, [_p]),
    "hgb_time_end": (_i, [_p, _p]),
    "hgb_flush_l2": (_i, [_p, _sz]),
    "hgb_set_profiling": (_i, [_p, _i]),
    "hgb_get_profile": (_i, [_p, _p, _p, _i]),
}

_lib = None


def lib():
    """Load libhgb200.so (once). Fails loudly if it was not built — there is no fallback implementation."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise HgbError("%s is missing: build it with `make -C harfang3d_b200` (or __graft_entry__.build()). "
                           "There is no CPU or PyTorch fallback for this path." % path)
        L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def _ptr(a, dtype, what):
    if a is None:
        return None
    if not isinstance(a, np.ndarray) or a.dtype != np.dtype(dtype) or not a.flags["C_CONTIGUOUS"]:
        raise TypeError("%s must be a C-contiguous numpy array of dtype %s" % (what, np.dtype(dtype)))
    return a.ctypes.data_as(C.c_void_p)


NOALLOC_MAX_ENTRIES = 1 << 24


def compact_visible_noalloc(device, d_vis, n, view_bit, d_out, out_capacity, d_count, stream=None):
    """hgb_compact_visible_noalloc: device POINTERS (ints) of the caller's buffers; enqueues on `stream` (cudaStream_t as int, None =
    default stream) and returns — synchronise before reading. Allocates nothing; n <= NOALLOC_MAX_ENTRIES."""
    L = lib()
    if L.hgb_compact_visible_noalloc(device, C.c_void_p(d_vis), n, view_bit, C.c_void_p(d_out), out_capacity, C.c_void_p(d_count), C.c_void_p(stream or 0)):
        raise HgbError(L.hgb_last_error().decode())


class Context:
    """One hgb_ctx = one hg::Scene (or one batch of scene instances)."""

    def __init__(self, device=0, use_graph=True, row_store=False, lookahead=0, tma_level=False, top_walk=0, overlap_sampler=False, slice_nodes=0, async_post=None, append_slack=0):
        """row_store=True: world rows are stored per thread (3 x 16 B at a 48-byte stride) instead of through the shared-memory
        transpose — an A/B switch for profiling (hgb_options.row_store)"""
        self._L = lib()
        # async_post: None = the library default (chosen per frame; HGB_ASYNC_POST=0/1/2 in the environment overrides), True / False = forced
        opt = (C.c_int * 9)(1 if use_graph else 0, 1 if row_store else 0, lookahead, 1 if tma_level else 0, (0 if async_post is None else (1 if async_post else -1)), top_walk, 1 if overlap_sampler else 0, slice_nodes, append_slack)  # lookahead: 0 default, -1 off, n positions
        self._h = self._L.hgb_create(device, C.cast(opt, C.c_void_p))
        if not self._h:
            raise HgbError(self._L.hgb_last_error().decode())
        self._h = C.c_void_p(self._h)
        self.n_slots = 0
        self.n_entries = [0, 0]
        self.n_bones = 0
        self.n_inst = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.hgb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise HgbError(self._L.hgb_last_error().decode())

    # ---- edit time ----
    def set_stream(self, cuda_stream_ptr):
        self._ck(self._L.hgb_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        self._ck(self._L.hgb_sync(self._h))

    def set_topology(self, parent_slot, used=None):
        n = len(parent_slot)
        self._ck(self._L.hgb_set_topology(self._h, n, _ptr(parent_slot, np.uint32, "parent_slot"), _ptr(used, np.uint8, "used")))
        self.n_slots = n

    def set_parents(self, slots, parent_slots):
        """incremental SetTransformParent; returns True when the layout had to be rebuilt (a level changed)"""
        rebuilt = C.c_int(0)
        self._ck(self._L.hgb_set_parents(self._h, len(slots), _ptr(slots, np.uint32, "slots"), _ptr(parent_slots, np.uint32, "parent_slots"), C.byref(rebuilt)))
        return bool(rebuilt.value)

    def append_nodes(self, slots, parent_slots, trs9):
        """new leaf transforms in free slots; returns True when the layout had to be rebuilt"""
        rebuilt = C.c_int(0)
        trs9 = np.ascontiguousarray(trs9, np.float32).reshape(-1, 9)
        assert trs9.shape[0] == len(slots)
        self._ck(self._L.hgb_append_nodes(self._h, len(slots), _ptr(slots, np.uint32, "slots"), _ptr(parent_slots, np.uint32, "parent_slots"),
                                          _ptr(trs9, np.float32, "trs9"), C.byref(rebuilt)))
        self.n_slots = max(self.n_slots, int(np.max(slots)) + 1) if len(slots) else self.n_slots
        return bool(rebuilt.value)

    def remove_nodes(self, slots):
        rebuilt = C.c_int(0)
        self._ck(self._L.hgb_remove_nodes(self._h, len(slots), _ptr(slots, np.uint32, "slots"), C.byref(rebuilt)))
        return bool(rebuilt.value)

    def append_display_entries(self, kind, mtx_slot, bounds_idx, flags=None, view=None):
        rebuilt = C.c_int(0)
        self._ck(self._L.hgb_append_display_entries(self._h, kind, len(mtx_slot), _ptr(mtx_slot, np.uint32, "mtx_slot"), _ptr(bounds_idx, np.uint32, "bounds_idx"),
                                                    _ptr(flags, np.uint8, "flags"), _ptr(view, np.uint16, "view"), C.byref(rebuilt)))
        self.n_entries[kind] += len(mtx_slot)
        return bool(rebuilt.value)

    def upload_trs(self, trs9, first=0):
        assert trs9.ndim == 2 and trs9.shape[1] == 9
        self._ck(self._L.hgb_upload_trs(self._h, first, trs9.shape[0], _ptr(trs9, np.float32, "trs9")))

    def set_preset_world(self, slots, mats12):
        self._ck(self._L.hgb_set_preset_world(self._h, len(slots), _ptr(slots, np.uint32, "slots"), _ptr(mats12, np.float32, "mats12")))

    def set_matrices(self, slots, mats12, local=False):
        """batched Scene::SetTransformWorldMatrix (local=False) / SetTransformLocalMatrix (local=True)"""
        self._ck(self._L.hgb_set_matrices(self._h, len(slots), _ptr(slots, np.uint32, "slots"), _ptr(mats12, np.float32, "mats12"), 1 if local else 0))

    def set_bounds(self, minmax6):
        self._ck(self._L.hgb_set_bounds(self._h, minmax6.shape[0], _ptr(minmax6, np.float32, "minmax6")))

    def set_display_entries(self, kind, mtx_slot, bounds_idx, flags=None, view=None):
        self._ck(self._L.hgb_set_display_entries(self._h, kind, len(mtx_slot), _ptr(mtx_slot, np.uint32, "mtx_slot"), _ptr(bounds_idx, np.uint32, "bounds_idx"),
                                                 _ptr(flags, np.uint8, "flags"), _ptr(view, np.uint16, "view")))
        self.n_entries[kind] = len(mtx_slot)

    def set_entry_flags(self, kind, first, flags):
        self._ck(self._L.hgb_set_entry_flags(self._h, kind, first, len(flags), _ptr(flags, np.uint8, "flags")))

    def set_skins(self, mtx_slot, bind_idx, bind_pose12):
        self._ck(self._L.hgb_set_skins(self._h, len(mtx_slot), _ptr(mtx_slot, np.uint32, "mtx_slot"), _ptr(bind_idx, np.uint32, "bind_idx"),
                                       bind_pose12.shape[0], _ptr(bind_pose12, np.float32, "bind_pose12")))
        self.n_bones = len(mtx_slot)

    def set_anim_keys(self, v3_t, v3_val, v3_tension, v3_bias, q_t, q_val):
        self._ck(self._L.hgb_set_anim_keys(self._h, len(v3_t), _ptr(v3_t, np.int64, "v3_t"), _ptr(v3_val, np.float32, "v3_val"),
                                           _ptr(v3_tension, np.float32, "v3_tension"), _ptr(v3_bias, np.float32, "v3_bias"), len(q_t),
                                           _ptr(q_t, np.int64, "q_t"), _ptr(q_val, np.float32, "q_val")))

    def set_tracks(self, tracks, n_play):
        self._ck(self._L.hgb_set_tracks(self._h, len(tracks), _ptr(tracks, TRACK_DTYPE, "tracks"), n_play))

    def set_enable_tracks(self, tracks, key_t, key_v):
        """"Enable" bool tracks (TRACK_DTYPE records, kind ignored) over the bool key pool (key_t int64, key_v uint8)"""
        self._ck(self._L.hgb_set_enable_tracks(self._h, len(tracks), _ptr(tracks, TRACK_DTYPE, "tracks"), len(key_t), _ptr(key_t, np.int64, "key_t"), _ptr(key_v, np.uint8, "key_v")))
        self._n_enable = len(np.unique(tracks["target_slot"])) if len(tracks) else 0

    def set_enable_state(self, slots, disabled):
        self._ck(self._L.hgb_set_enable_state(self._h, len(slots), _ptr(slots, np.uint32, "slots"), _ptr(disabled, np.uint8, "disabled")))

    def get_enable_state(self):
        """-> (target slots ascending, disabled 0/1) of the enable-tracked nodes"""
        n = max(getattr(self, "_n_enable", 0), 1)
        slots, dis, cnt = np.zeros(n, np.uint32), np.zeros(n, np.uint8), C.c_uint32(0)
        self._ck(self._L.hgb_get_enable_state(self._h, n, _ptr(slots, np.uint32, "slots"), _ptr(dis, np.uint8, "dis"), C.byref(cnt)))
        return slots[: cnt.value], dis[: cnt.value]

    def set_views(self, frustum24):
        f = np.ascontiguousarray(frustum24, np.float32).reshape(-1, 24)
        self._ck(self._L.hgb_set_views(self._h, f.shape[0], _ptr(f, np.float32, "frustum24")))

    def set_instance_views(self, frustum24, inst_entry_start):
        f = np.ascontiguousarray(frustum24, np.float32).reshape(-1, 24)
        self._ck(self._L.hgb_set_instance_views(self._h, f.shape[0], _ptr(f, np.float32, "frustum24"), _ptr(inst_entry_start, np.uint32, "inst_entry_start")))
        self.n_inst = f.shape[0]

    # ---- per frame ----
    def update(self, play_t=None, fused_cull=False):
        n = 0 if play_t is None else len(play_t)
        self._ck(self._L.hgb_update(self._h, n, _ptr(play_t, np.int64, "play_t"), 1 if fused_cull else 0))

    def cull(self):
        self._ck(self._L.hgb_cull(self._h))

    def cull_list(self, mtx_slot, bounds_idx, bounds6, frustum24):
        """CullModelDisplayLists on a caller-owned list -> surviving entry indices, in order"""
        out, n = np.empty(max(len(mtx_slot), 1), np.uint32), C.c_uint32(0)
        f = np.ascontiguousarray(frustum24, np.float32).reshape(24)
        self._ck(self._L.hgb_cull_list(self._h, len(mtx_slot), _ptr(mtx_slot, np.uint32, "mtx_slot"), _ptr(bounds_idx, np.uint32, "bounds_idx"),
                                       bounds6.shape[0], _ptr(bounds6, np.float32, "bounds6"), _ptr(f, np.float32, "frustum24"),
                                       _ptr(out, np.uint32, "out"), C.byref(n)))
        return out[: n.value]

    def sort_keys(self, kind, view, view_mat12):
        """ComputeModelDisplayListSortKeys for the survivors of list (kind, view) -> uint32 keys in visible-list order"""
        m = np.ascontiguousarray(view_mat12, np.float32).reshape(12)
        self._ck(self._L.hgb_sort_keys(self._h, kind, view, _ptr(m, np.float32, "view_mat12")))
        out, n = np.empty(max(self.n_entries[kind], 1), np.uint32), C.c_uint32(0)
        self._ck(self._L.hgb_get_sort_keys(self._h, kind, view, _ptr(out, np.uint32, "out"), len(out), C.byref(n)))
        return out[: n.value]

    def set_objects(self, mtx_slot, obounds_idx, obounds6, seg=None, n_seg=1):
        """object-carrying nodes for Scene::GetMinMax / GetNodesMinMax (one segment per scene instance)"""
        self.n_seg = n_seg
        self._ck(self._L.hgb_set_objects(self._h, len(mtx_slot), _ptr(mtx_slot, np.uint32, "mtx_slot"), _ptr(obounds_idx, np.uint32, "obounds_idx"),
                                         obounds6.shape[0], _ptr(obounds6, np.float32, "obounds6"), _ptr(seg, np.uint32, "seg"), n_seg))

    def minmax(self, fetch=True):
        """-> (minmax[n_seg, 6], contributing node count[n_seg])"""
        self._ck(self._L.hgb_minmax(self._h))
        if not fetch:
            return None
        out, cnt = np.zeros((self.n_seg, 6), np.float32), np.zeros(self.n_seg, np.uint32)
        self._ck(self._L.hgb_get_minmax(self._h, _ptr(out, np.float32, "out"), _ptr(cnt, np.uint32, "cnt")))
        return out, cnt

    # ---- multi-GPU result gather over peer memory ----
    def peer_init(self, rank, world, dst, entries_per_rank, aux_words_per_rank=None):
        """-> (inbox handle, mailbox handle), 64 opaque bytes each, to be exchanged between the ranks (see peer_setup)"""
        cap = np.ascontiguousarray(entries_per_rank, np.uint32)
        aux = None if aux_words_per_rank is None else np.ascontiguousarray(aux_words_per_rank, np.uint32)
        inbox, mail = C.create_string_buffer(64), C.create_string_buffer(64)
        self._ck(self._L.hgb_peer_init(self._h, rank, world, dst, _ptr(cap, np.uint32, "cap"), _ptr(aux, np.uint32, "aux"), C.addressof(inbox), C.addressof(mail)))
        self._peer = (rank, world, dst, int(cap.sum()), 0 if aux is None else int(aux.sum()))
        return inbox.raw, mail.raw

    def peer_connect(self, inbox_of_dst, mailboxes):
        blob = C.create_string_buffer(b"".join(mailboxes), 64 * len(mailboxes))
        dst = C.create_string_buffer(inbox_of_dst, 64)
        self._ck(self._L.hgb_peer_connect(self._h, C.addressof(dst), C.addressof(blob)))

    def peer_close(self):
        self._ck(self._L.hgb_peer_close(self._h))
        self._peer = None

    def peer_setup(self, dist, entries_per_rank, aux_words_per_rank=None, dst=0, group=None):
        """init + handle exchange over torch.distributed + connect. Collective: every rank calls it, and every rank leaves it the same
        way — a failure on ANY rank (no IPC on the box, no peer access between two GPUs) closes the gather on all of them and raises
        HgbError everywhere, so the callers can take the same decision (bench.py: the NCCL exchange instead)."""
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        inbox = mail = err = None
        try:
            inbox, mail = self.peer_init(rank, world, dst, entries_per_rank, aux_words_per_rank)
        except HgbError as e:
            err = "rank %d, init: %s" % (rank, e)
        handles = [None] * world
        dist.all_gather_object(handles, (inbox, mail, err), group=group)
        if not any(h[2] for h in handles):
            try:
                self.peer_connect(handles[dst][0], [h[1] for h in handles])
            except HgbError as e:
                err = "rank %d, connect: %s" % (rank, e)
        errs = [None] * world
        dist.all_gather_object(errs, err or handles[rank][2], group=group)
        failed = [e for e in errs if e] + [h[2] for h in handles if h[2]]
        if failed:
            self.peer_close()
            raise HgbError("peer gather unavailable: " + failed[0])
        dist.barrier(group)

    def peer_status(self):
        step, err = C.c_uint32(0), C.c_uint32(0)
        self._ck(self._L.hgb_peer_status(self._h, C.byref(step), C.byref(err)))
        return step.value, err.value

    def peer_result(self, with_idx=True):
        """destination rank: (rank offsets[world+1], concatenated global entry indices, global per-instance offsets + total)"""
        rank, world, dst, cap, n_aux = self._peer
        off = np.zeros(world + 1, np.uint32)
        idx = np.zeros(max(cap, 1), np.uint32) if with_idx else None
        aux = np.zeros(n_aux + 1, np.uint32)
        self._ck(self._L.hgb_peer_result(self._h, _ptr(off, np.uint32, "off"), _ptr(idx, np.uint32, "idx"), 0 if idx is None else len(idx), _ptr(aux, np.uint32, "aux"), len(aux)))
        return off, (idx[: off[-1]] if with_idx else None), aux

    def bone_matrices(self, with_previous=False):
        self._ck(self._L.hgb_bone_matrices(self._h, 1 if with_previous else 0))

    # ---- results ----
    def get_worlds(self, first=0, count=None, previous=False, out=None):
        count = self.n_slots - first if count is None else count
        out = np.empty((count, 12), np.float32) if out is None else out
        self._ck(self._L.hgb_get_worlds(self._h, first, count, _ptr(out, np.float32, "out"), 1 if previous else 0))
        return out

    def get_trs(self, first=0, count=None):
        count = self.n_slots - first if count is None else count
        out = np.empty((count, 9), np.float32)
        self._ck(self._L.hgb_get_trs(self._h, first, count, _ptr(out, np.float32, "out")))
        return out

    def get_visible_count(self, kind=0, view=0):
        n = C.c_uint32(0)
        self._ck(self._L.hgb_get_visible_count(self._h, kind, view, C.byref(n)))
        return n.value

    def get_visible(self, kind=0, view=0, out=None):
        out = np.empty(max(self.n_entries[kind], 1), np.uint32) if out is None else out
        n = C.c_uint32(0)
        self._ck(self._L.hgb_get_visible(self._h, kind, view, _ptr(out, np.uint32, "out"), len(out), C.byref(n)))
        return out[: n.value]

    def cull_skinned_list(self, bone_off, mtx_slot, bind_idx, bind_pose12, bounds_idx, bounds6, frustum24):
        """opt-in cull of skinned display lists (union of the per-bone boxes) -> surviving list indices, in order"""
        n = len(bone_off) - 1
        out, cnt = np.empty(max(n, 1), np.uint32), C.c_uint32(0)
        f = np.ascontiguousarray(frustum24, np.float32).reshape(24)
        self._ck(self._L.hgb_cull_skinned_list(self._h, n, _ptr(bone_off, np.uint32, "bone_off"), _ptr(mtx_slot, np.uint32, "mtx_slot"), _ptr(bind_idx, np.uint32, "bind_idx"),
                                               bind_pose12.shape[0], _ptr(bind_pose12, np.float32, "bind_pose12"), _ptr(bounds_idx, np.uint32, "bounds_idx"),
                                               bounds6.shape[0], _ptr(bounds6, np.float32, "bounds6"), _ptr(f, np.float32, "frustum24"), _ptr(out, np.uint32, "out"),
                                               C.byref(cnt)))
        return out[: cnt.value]

    def get_visible_ptr(self, kind=0, view=0):
        """the visible list in the context's own pinned buffer (valid until the next call for this kind): one round trip"""
        ptr, n = C.POINTER(C.c_uint32)(), C.c_uint32(0)
        self._ck(self._L.hgb_get_visible_ptr(self._h, kind, view, C.byref(ptr), C.byref(n)))
        return np.ctypeslib.as_array(ptr, shape=(n.value,)) if n.value else np.zeros(0, np.uint32)

    def set_entry_records(self, kind, rec16, first=0):
        """opaque 16-byte records of entries [first, first + n) of `kind` ((n, 16) uint8 or (n, 4) uint32)"""
        r = np.ascontiguousarray(rec16).view(np.uint8).reshape(-1, 16)
        self._ck(self._L.hgb_set_entry_records(self._h, kind, first, r.shape[0], r.ctypes.data_as(C.c_void_p)))

    def get_visible_records(self, kind=0, view=0):
        """the visible list as the entries' 16-byte records (device-side gather) -> (n, 16) uint8 view of the context's pinned buffer"""
        ptr, n = C.c_void_p(), C.c_uint32(0)
        self._ck(self._L.hgb_get_visible_records_ptr(self._h, kind, view, C.byref(ptr), C.byref(n)))
        if not n.value:
            return np.zeros((0, 16), np.uint8)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n.value * 16,)).reshape(-1, 16)

    def get_instance_offsets(self):
        out = np.empty(self.n_inst + 1, np.uint32)
        self._ck(self._L.hgb_get_instance_offsets(self._h, _ptr(out, np.uint32, "out")))
        return out

    def get_bone_matrices(self, previous=False):
        out = np.empty((self.n_bones, 16), np.float32)
        self._ck(self._L.hgb_get_bone_matrices(self._h, _ptr(out, np.float32, "out"), 1 if previous else 0))
        return out

    # ---- introspection ----
    def device_ptr(self, which):
        return self._L.hgb_device_ptr(self._h, which)

    @property
    def launch_count(self):
        return self._L.hgb_launch_count(self._h)

    @property
    def frame_was_async(self):
        return bool(self._L.hgb_frame_was_async(self._h))

    @property
    def rebuild_count(self):
        return self._L.hgb_rebuild_count(self._h)

    @property
    def level_count(self):
        return self._L.hgb_level_count(self._h)

    @property
    def identity_order(self):
        return bool(self._L.hgb_is_identity_order(self._h))

    def time_begin(self):
        self._ck(self._L.hgb_time_begin(self._h))

    def time_end(self):
        ms = C.c_float(0)
        self._ck(self._L.hgb_time_end(self._h, C.byref(ms)))
        return ms.value

    def set_profiling(self, on):
        self._ck(self._L.hgb_set_profiling(self._h, 1 if on else 0))

    def get_profile(self, reset=True):
        """{kernel class: (total ms, launches)} accumulated since the last reset"""
        ms, n = np.zeros(len(KERNEL_CLASSES), np.float64), np.zeros(len(KERNEL_CLASSES), np.uint64)
        self._ck(self._L.hgb_get_profile(self._h, _ptr(ms, np.float64, "ms"), _ptr(n, np.uint64, "n"), 1 if reset else 0))
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(KERNEL_CLASSES)}

    def flush_l2(self, nbytes=256 << 20):
        self._ck(self._L.hgb_flush_l2(self._h, nbytes))
