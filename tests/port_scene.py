"""TEST INFRASTRUCTURE: the same frame as scene_io.RefScene, computed with the plain-C port (oracle/hg_port.c). Used where the
compiled reference (oracle/_ref) is unavailable, and to cross-check the two oracles against each other."""
import numpy as np

from harfang3d_b200 import hostmirror
from tests.oracle_lib import P


class PortScene:
    def __init__(self, port, fs, tables=None):
        self.L, self.fs = port, fs
        self.t = hostmirror.build_tables(fs) if tables is None else tables
        self.trs_now = np.ascontiguousarray(fs.trs, np.float32).copy()
        self.clocks = hostmirror.PlayClocks(fs)
        self.w = np.zeros((fs.n, 12), np.float32)
        self.w_prev = np.zeros((fs.n, 12), np.float32)

    def close(self):
        pass

    def update(self, dt):
        fs, L = self.fs, self.L
        self.w, self.w_prev = self.w_prev, self.w
        times = self.clocks.step(int(dt))
        out3, out4 = np.zeros(3, np.float32), np.zeros(4, np.float32)
        for rec in self.t["tracks"]:
            at = int(times[rec["play"]])
            if at == hostmirror.NOT_PLAYING:
                continue
            off, cnt, kind, node = int(rec["key_off"]), int(rec["key_count"]), int(rec["kind"]), int(rec["target_slot"])
            if kind == 3:
                if L.hgp_eval_quat_track(cnt, P(fs.q_t[off:off + cnt]), P(np.ascontiguousarray(fs.q_val[off:off + cnt])), at, P(out4)):
                    L.hgp_quat_to_euler(P(out4), P(out3))
                    self.trs_now[node, 3:6] = out3
            else:
                out3[:] = self.trs_now[node, kind * 3:kind * 3 + 3]
                if L.hgp_eval_vec3_track(cnt, P(fs.v3_t[off:off + cnt]), P(np.ascontiguousarray(fs.v3_val[off:off + cnt])), P(fs.v3_tension[off:off + cnt]),
                                         P(fs.v3_bias[off:off + cnt]), at, P(out3)):
                    self.trs_now[node, kind * 3:kind * 3 + 3] = out3
        L.hgp_compute_world_matrices(fs.n, P(self.trs_now), P(fs.parent), None, None, P(self.w))

    def worlds(self, previous=False):
        return self.w_prev if previous else self.w

    def trs(self, nodes):
        return self.trs_now[np.asarray(nodes)]

    def cull(self, kind, frustum24):
        """returns (positions in the reference's un-culled list, mtx_idx) like RefScene.cull"""
        t = self.t
        flags = t["ent%d_flags" % kind]
        enabled = np.nonzero(flags == 0)[0]
        mtx = np.ascontiguousarray(t["ent%d_mtx_slot" % kind][enabled])
        bidx = np.ascontiguousarray(t["ent%d_bounds_idx" % kind][enabled])
        vis = np.zeros(max(len(mtx), 1), np.uint32)
        k = self.L.hgp_cull(len(mtx), P(mtx), P(bidx), P(t["bounds"]), P(self.w), P(np.ascontiguousarray(frustum24, np.float32)), P(vis))
        return vis[:k].copy(), mtx[vis[:k]].copy()

    def bone_matrices(self, previous=False, kind=2):
        t = self.t
        n = len(t["skin_mtx_slot"])
        out = np.zeros((n, 16), np.float32)
        if n:
            self.L.hgp_bone_matrices(n, P(t["skin_mtx_slot"]), P(t["skin_bind_idx"]), P(self.w_prev if previous else self.w), P(t["bind_pose"]), P(out))
        return out
