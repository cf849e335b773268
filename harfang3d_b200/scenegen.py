"""Seeded synthetic scenes for the BASELINE.json configs, as flat numpy arrays ("FlatScene").

A FlatScene is a neutral description: tests feed the SAME arrays to the reference (through the flat binding csrc/facade/flat/hg_flat.cpp, using the
reference's public Scene API) and to the product (through hostmirror.upload -> C ABI). Node i owns transform slot i, which is how
the reference numbers them when nodes are created in order (vector_list LIFO free list, foundation/vector_list.h:182-218).
Workload definitions follow SURVEY.md §8(d).
"""
from dataclasses import dataclass, field

import numpy as np

NS = 1_000_000_000
FRAME_DT = 16_666_667  # ns, 60 Hz
KEY_DT = 50_000_000    # ns, 20 fps key spacing (SceneAnim::frame_duration default, engine/scene.h:116)


@dataclass
class FlatScene:
    trs: np.ndarray                      # (n, 9) f32: pos, rot (Euler YXZ, radians), scale
    parent: np.ndarray                   # (n,) i32 node index or -1
    disabled: np.ndarray                 # (n,) u8
    model: np.ndarray                    # (n,) i32 model index or -1 (no object component)
    n_mats: np.ndarray                   # (n,) u8 material slots of the object
    transparent_mask: np.ndarray         # (n,) u32 bit k -> material k is alpha blended
    models: list = field(default_factory=list)   # dicts: bounds (L,6) f32, mats (L,) u16, bone_tables [arrays u16] | None, bind_pose (B,12) | None
    bones: dict = field(default_factory=dict)    # object node -> (B,) i32 bone node indices (-1 = invalid ref)
    # animation (flat): one hg::Anim per entry of anim_node; tracks reference their anim
    anim_node: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    anim_use_quat: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint8))
    trk_anim: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    trk_kind: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint8))      # 0 pos 1 rot(euler) 2 scale 3 rot(quat)
    trk_key_off: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint32))
    trk_key_count: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint32))
    v3_t: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int64))
    v3_val: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    v3_tension: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    v3_bias: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    q_t: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int64))
    q_val: np.ndarray = field(default_factory=lambda: np.zeros((0, 4), np.float32))
    scene_anims: list = field(default_factory=list)  # (anim index array i32, t_start, t_end)
    plays: list = field(default_factory=list)        # (scene_anim index, loop_mode 0 once/1 infinite/2 loop, t_scale float)
    camera: dict = field(default_factory=dict)       # znear, zfar, fov, ar_x, ar_y, world12 (camera world matrix)
    # batches of independent scenes laid out back to back in one slot space
    instance_start: np.ndarray = None                # (n_inst+1,) node offsets, or None for a single scene
    instance_cameras: np.ndarray = None              # (n_inst, 12) camera world matrix per instance

    @property
    def n(self):
        return self.trs.shape[0]


def level_sizes(n_nodes, depth, fanout):
    """1, f, f^2, ... capped, last level topped up so the total is exactly n_nodes"""
    sizes, total = [], 0
    for d in range(depth - 1):
        s = min(fanout ** d, n_nodes - total - (depth - 1 - d))
        sizes.append(int(s))
        total += int(s)
    sizes.append(int(n_nodes - total))
    assert all(s > 0 for s in sizes) and sum(sizes) == n_nodes
    return sizes


def tree_parents(sizes):
    """BFS-numbered tree: children of a level are spread evenly over the previous level, siblings adjacent"""
    n = sum(sizes)
    parent = np.full(n, -1, np.int32)
    start = 0
    for d in range(1, len(sizes)):
        prev_start, prev, cur = start, sizes[d - 1], sizes[d]
        start += prev
        j = np.arange(cur, dtype=np.int64)
        parent[start:start + cur] = (prev_start + (j * prev) // cur).astype(np.int32)
    return parent


def random_trs(rng, n, parent, root_extent=50.0, child_extent=5.0):
    trs = np.empty((n, 9), np.float32)
    trs[:, 0:3] = rng.uniform(-child_extent, child_extent, (n, 3))
    roots = parent < 0
    trs[roots, 0:3] = rng.uniform(-root_extent, root_extent, (int(roots.sum()), 3))
    trs[:, 3:6] = rng.uniform(-np.pi, np.pi, (n, 3))
    trs[:, 6:9] = rng.uniform(0.5, 2.0, (n, 3))
    return trs


def box_models(rng, count):
    models = []
    for m in range(count):
        if m == 0:
            b = np.array([[-0.5, -0.5, -0.5, 0.5, 0.5, 0.5]], np.float32)  # MinMaxFromPositionSize({0,0,0},{1,1,1})
        else:
            lo = rng.uniform(-2, 0, 3)
            b = np.concatenate([lo, lo + rng.uniform(0.2, 3, 3)]).astype(np.float32)[None]
        models.append(dict(bounds=b, mats=np.zeros(1, np.uint16), bone_tables=None, bind_pose=None))
    return models


IDENTITY12 = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
DEFAULT_CAMERA = dict(znear=0.1, zfar=500.0, fov=float(np.float32(np.pi / 3)), ar_x=float(np.float32(16 / 9)), ar_y=1.0, world12=IDENTITY12.copy())


def random_camera_worlds(rng, n, extent=20.0):
    """camera world matrices (n,12) f32: random rotation (computed in f64, rounded once) + translation, unit scale"""
    q = rng.normal(0, 1, (n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    x, y, z, w = q.T
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                  2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                  2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1).reshape(n, 3, 3)
    m = np.zeros((n, 3, 4))
    m[:, :, :3] = R
    m[:, :, 3] = rng.uniform(-extent, extent, (n, 3))
    return m.reshape(n, 12).astype(np.float32)


def static_scene(n_nodes, depth, fanout, n_models=1, seed=0, disabled_frac=0.01, scale_world=1.0):
    """C1 (10k, depth 4, fan-out 21) and C2 (1M, depth 8, fan-out 7): every node has a transform and an object."""
    rng = np.random.default_rng(seed)
    parent = tree_parents(level_sizes(n_nodes, depth, fanout))
    trs = random_trs(rng, n_nodes, parent)
    trs[:, 6:9] = rng.uniform(0.8, 1.25, (n_nodes, 3)) * scale_world  # keep deep chains from exploding
    trs[parent < 0, 0:3] = rng.uniform(-2, 2, (int((parent < 0).sum()), 3))  # the camera (at the origin) sits inside the scene: a cone of it is visible
    return FlatScene(trs=trs, parent=parent, disabled=(rng.random(n_nodes) < disabled_frac).astype(np.uint8),
                     model=rng.integers(0, n_models, n_nodes).astype(np.int32), n_mats=np.ones(n_nodes, np.uint8),
                     transparent_mask=np.zeros(n_nodes, np.uint32), models=box_models(rng, n_models), camera=dict(DEFAULT_CAMERA))


def config_c1(seed=1):
    return static_scene(10_000, 4, 21, n_models=1, seed=seed)


def config_c2(seed=2, n_nodes=1_000_000):
    return static_scene(n_nodes, 8, 7, n_models=16, seed=seed)


def add_position_tracks(fs, nodes, rng, n_keys=32, play_loop=True):
    """one Anim per node with a single Hermite "Position" track, all driven by ONE SceneAnim played in a loop (C4's animated part)"""
    a = len(nodes)
    base_anim, base_trk, base_key = len(fs.anim_node), len(fs.trk_anim), len(fs.v3_t)
    fs.anim_node = np.concatenate([fs.anim_node, nodes.astype(np.int32)])
    fs.anim_use_quat = np.concatenate([fs.anim_use_quat, np.zeros(a, np.uint8)])
    fs.trk_anim = np.concatenate([fs.trk_anim, base_anim + np.arange(a, dtype=np.int32)])
    fs.trk_kind = np.concatenate([fs.trk_kind, np.zeros(a, np.uint8)])
    fs.trk_key_off = np.concatenate([fs.trk_key_off, (base_key + np.arange(a, dtype=np.int64) * n_keys).astype(np.uint32)])
    fs.trk_key_count = np.concatenate([fs.trk_key_count, np.full(a, n_keys, np.uint32)])
    t = np.tile(np.arange(n_keys, dtype=np.int64) * KEY_DT, a)
    fs.v3_t = np.concatenate([fs.v3_t, t])
    centre = fs.trs[nodes, 0:3]
    val = rng.random((a * n_keys, 3), dtype=np.float32)  # f32 draws: the 16M-node batch has 10^8 keys
    val *= np.float32(4)
    val -= np.float32(2)
    val += np.repeat(centre, n_keys, axis=0)
    fs.v3_val = np.concatenate([fs.v3_val, val]) if len(fs.v3_val) else val
    fs.v3_tension = np.concatenate([fs.v3_tension, rng.random(a * n_keys, dtype=np.float32) - np.float32(0.5)])
    fs.v3_bias = np.concatenate([fs.v3_bias, rng.random(a * n_keys, dtype=np.float32) - np.float32(0.5)])
    fs.scene_anims.append((base_anim + np.arange(a, dtype=np.int32), 0, (n_keys - 1) * KEY_DT))
    fs.plays.append((len(fs.scene_anims) - 1, 2 if play_loop else 0, 1.0))
    return base_trk


def concat_scenes(scenes):
    """lay independent scenes back to back in one slot space (batch of scene instances)"""
    off = np.cumsum([0] + [s.n for s in scenes])
    first = scenes[0]
    assert all(len(s.anim_node) == 0 for s in scenes), "concat before adding animation"
    parent = np.concatenate([np.where(s.parent >= 0, s.parent + o, -1) for s, o in zip(scenes, off)]).astype(np.int32)
    fs = FlatScene(trs=np.concatenate([s.trs for s in scenes]), parent=parent, disabled=np.concatenate([s.disabled for s in scenes]),
                   model=np.concatenate([s.model for s in scenes]), n_mats=np.concatenate([s.n_mats for s in scenes]),
                   transparent_mask=np.concatenate([s.transparent_mask for s in scenes]), models=first.models, camera=dict(first.camera))
    fs.instance_start = off.astype(np.uint32)
    return fs


def config_c4(seed=4, n_scenes=16, nodes_per_scene=1_000_000, animated_frac=0.25, n_keys=32):
    """16 scenes x 1M nodes (each respects the reference's 2^24-1 slot cap), depth 8, 25 % of the nodes with a Position track"""
    rng = np.random.default_rng(seed)
    base = static_scene(nodes_per_scene, 8, 7, n_models=16, seed=seed)
    scenes = []
    for s in range(n_scenes):
        sc = FlatScene(trs=base.trs.copy(), parent=base.parent, disabled=base.disabled, model=base.model, n_mats=base.n_mats,
                       transparent_mask=base.transparent_mask, models=base.models, camera=base.camera)
        sc.trs[:, 0:3] += rng.uniform(-0.5, 0.5, 3).astype(np.float32)  # scenes differ
        scenes.append(sc)
    fs = concat_scenes(scenes)
    n_anim = int(fs.n * animated_frac)
    nodes = np.sort(rng.choice(fs.n, n_anim, replace=False)).astype(np.int32)
    add_position_tracks(fs, nodes, rng, n_keys=n_keys)
    fs.instance_start = None  # one shared camera: the batch is culled as one list
    return fs


def config_c3(seed=3, n_chars=4096, n_bones=64, n_keys=32):
    """Skinned characters: per character 1 skinned object node + n_bones bone nodes (tree, total depth 8) + 1 rigid prop parented to the
    last bone (so the cull has something to do: the reference never culls skinned lists, scene_forward_pipeline.cpp:560-562).
    Every bone has Position/Rotation/Scale tracks; odd characters rotate with quaternion tracks (AF_UseQuaternionForRotation)."""
    rng = np.random.default_rng(seed)
    per = n_bones + 2
    bone_parent = tree_parents(level_sizes(n_bones, 7, 2))  # local to the bone block, -1 = child of the object node
    n = n_chars * per
    parent = np.full(n, -1, np.int32)
    model = np.full(n, -1, np.int32)
    for c in range(n_chars):
        b = c * per
        parent[b + 1:b + 1 + n_bones] = np.where(bone_parent >= 0, bone_parent + b + 1, b)
        parent[b + 1 + n_bones] = b + n_bones  # prop on the last bone
        model[b] = 0
        model[b + 1 + n_bones] = 1
    trs = random_trs(rng, n, parent, child_extent=1.0)
    trs[:, 6:9] = rng.uniform(0.9, 1.1, (n, 3))
    roots = np.arange(n_chars) * per  # a crowd in front of and around the camera: part of it is inside the frustum
    trs[roots, 0:3] = np.stack([rng.uniform(-60, 60, n_chars), rng.uniform(-5, 5, n_chars), rng.uniform(-20, 100, n_chars)], axis=1)
    half = n_bones // 2
    bind = np.zeros((n_bones, 12), np.float32)
    bind[:, [0, 5, 10]] = 1
    bind[:, [3, 7, 11]] = rng.uniform(-1, 1, (n_bones, 3))
    skinned = dict(bounds=np.array([[-1, -1, -1, 1, 1, 1], [-1, 0, -1, 1, 2, 1]], np.float32), mats=np.zeros(2, np.uint16),
                   bone_tables=[np.arange(0, half, dtype=np.uint16), np.arange(half, n_bones, dtype=np.uint16)], bind_pose=bind)
    prop = dict(bounds=np.array([[-0.2, -0.2, -0.2, 0.2, 0.2, 0.2]], np.float32), mats=np.zeros(1, np.uint16), bone_tables=None, bind_pose=None)
    fs = FlatScene(trs=trs, parent=parent, disabled=np.zeros(n, np.uint8), model=model, n_mats=np.ones(n, np.uint8),
                   transparent_mask=np.zeros(n, np.uint32), models=[skinned, prop], camera=dict(DEFAULT_CAMERA))
    bone_nodes = (np.arange(n_chars, dtype=np.int64)[:, None] * per + 1 + np.arange(n_bones)[None, :]).astype(np.int32)
    for c in range(n_chars):
        fs.bones[c * per] = bone_nodes[c].copy()
    # animation: one Anim per bone, 3 tracks each
    a = n_chars * n_bones
    fs.anim_node = bone_nodes.reshape(-1).copy()
    use_quat = (np.arange(n_chars) % 2 == 1)
    fs.anim_use_quat = np.repeat(use_quat, n_bones).astype(np.uint8)
    t_keys = np.arange(n_keys, dtype=np.int64) * KEY_DT
    kinds = np.tile(np.array([0, 1, 2], np.uint8), a)
    fs.trk_anim = np.repeat(np.arange(a, dtype=np.int32), 3)
    quat_anim = np.repeat(fs.anim_use_quat.astype(bool), 3) & (kinds == 1)
    kinds[quat_anim] = 3
    fs.trk_kind = kinds
    fs.trk_key_count = np.full(3 * a, n_keys, np.uint32)
    is_q = kinds == 3
    off = np.zeros(3 * a, np.int64)
    off[~is_q] = np.arange(int((~is_q).sum()), dtype=np.int64) * n_keys
    off[is_q] = np.arange(int(is_q.sum()), dtype=np.int64) * n_keys
    fs.trk_key_off = off.astype(np.uint32)
    n3, nq = int((~is_q).sum()), int(is_q.sum())
    fs.v3_t = np.tile(t_keys, n3)
    v3_kind = np.repeat(kinds[~is_q], n_keys)
    v3_node = np.repeat(fs.anim_node[fs.trk_anim[~is_q]], n_keys)
    val = np.empty((n3 * n_keys, 3), np.float32)
    m0, m1, m2 = v3_kind == 0, v3_kind == 1, v3_kind == 2
    val[m0] = trs[v3_node[m0], 0:3] + rng.uniform(-0.3, 0.3, (int(m0.sum()), 3))
    val[m1] = rng.uniform(-np.pi, np.pi, (int(m1.sum()), 3))
    val[m2] = rng.uniform(0.9, 1.1, (int(m2.sum()), 3))
    fs.v3_val = val
    fs.v3_tension = rng.uniform(-0.5, 0.5, n3 * n_keys).astype(np.float32)
    fs.v3_bias = rng.uniform(-0.5, 0.5, n3 * n_keys).astype(np.float32)
    fs.q_t = np.tile(t_keys, nq)
    fs.q_val = rng.normal(0, 1, (nq * n_keys, 4)).astype(np.float32)
    for c in range(n_chars):
        fs.scene_anims.append((np.arange(c * n_bones, (c + 1) * n_bones, dtype=np.int32), 0, int(t_keys[-1])))
        fs.plays.append((c, 2, 1.0))
    return fs


def config_c5(seed=5, n_instances=8192, nodes_per_instance=2048, depth=6):
    """RL-style batch: independent instances, one camera each"""
    rng = np.random.default_rng(seed)
    fan = max(2, int(round(nodes_per_instance ** (1.0 / (depth - 1)))))
    sizes = level_sizes(nodes_per_instance, depth, fan)
    parent1 = tree_parents(sizes)
    n = n_instances * nodes_per_instance
    off = (np.arange(n_instances + 1, dtype=np.int64) * nodes_per_instance)
    parent = np.tile(parent1, n_instances).astype(np.int64)
    parent = np.where(parent >= 0, parent + np.repeat(off[:-1], nodes_per_instance), -1).astype(np.int32)
    trs = random_trs(rng, n, parent, root_extent=2.0)  # every instance lives around its own origin; its camera is inside it
    trs[:, 6:9] = rng.uniform(0.8, 1.25, (n, 3))
    fs = FlatScene(trs=trs, parent=parent, disabled=(rng.random(n) < 0.01).astype(np.uint8), model=rng.integers(0, 4, n).astype(np.int32),
                   n_mats=np.ones(n, np.uint8), transparent_mask=np.zeros(n, np.uint32), models=box_models(rng, 4), camera=dict(DEFAULT_CAMERA))
    fs.instance_start = off.astype(np.uint32)
    fs.instance_cameras = random_camera_worlds(rng, n_instances, extent=8.0)
    return fs


def slice_instances(fs, i0, i1):
    """The sub-batch [i0, i1) of an instance batch without animation (config_c5 style): what one rank owns when the batch is sharded
    across GPUs (sharding.partition_instances). Node indices are rebased to the slice; models are shared (replicated per rank)."""
    assert fs.instance_start is not None and len(fs.anim_node) == 0 and not fs.bones
    lo, hi = int(fs.instance_start[i0]), int(fs.instance_start[i1])
    parent = fs.parent[lo:hi].astype(np.int64)
    parent = np.where(parent >= 0, parent - lo, -1).astype(np.int32)
    sub = FlatScene(trs=fs.trs[lo:hi].copy(), parent=parent, disabled=fs.disabled[lo:hi].copy(), model=fs.model[lo:hi].copy(),
                    n_mats=fs.n_mats[lo:hi].copy(), transparent_mask=fs.transparent_mask[lo:hi].copy(), models=fs.models, camera=dict(fs.camera))
    sub.instance_start = (fs.instance_start[i0:i1 + 1].astype(np.int64) - lo).astype(np.uint32)
    sub.instance_cameras = fs.instance_cameras[i0:i1].copy()
    return sub


def extract_range(fs, lo, hi):
    """Nodes [lo, hi) of a batch as a stand-alone FlatScene (the range must be closed under parents — e.g. one scene of config_c4 —
    and must not contain skins). Animation is carried over: the Anims bound to nodes of the range, with their tracks and keys, and
    every SceneAnim / play restricted to them (same clocks). Used by the full-size parity tests: the reference runs one scene of the
    batch while the GPU runs all of them."""
    assert not fs.bones
    parent = fs.parent[lo:hi].astype(np.int64)
    assert ((parent < 0) | ((parent >= lo) & (parent < hi))).all(), "range is not closed under parents"
    sub = FlatScene(trs=fs.trs[lo:hi].copy(), parent=np.where(parent >= 0, parent - lo, -1).astype(np.int32), disabled=fs.disabled[lo:hi].copy(),
                    model=fs.model[lo:hi].copy(), n_mats=fs.n_mats[lo:hi].copy(), transparent_mask=fs.transparent_mask[lo:hi].copy(), models=fs.models,
                    camera=dict(fs.camera))
    if len(fs.anim_node):
        keep_anim = (fs.anim_node >= lo) & (fs.anim_node < hi)
        new_anim = np.cumsum(keep_anim) - 1
        sub.anim_node = (fs.anim_node[keep_anim] - lo).astype(np.int32)
        sub.anim_use_quat = fs.anim_use_quat[keep_anim].copy()
        keep_trk = keep_anim[fs.trk_anim]
        sub.trk_anim = new_anim[fs.trk_anim[keep_trk]].astype(np.int32)
        sub.trk_kind = fs.trk_kind[keep_trk].copy()
        sub.trk_key_count = fs.trk_key_count[keep_trk].copy()
        is_q = sub.trk_kind == 3
        cnt = sub.trk_key_count.astype(np.int64)
        off = np.zeros(len(cnt), np.int64)
        for m in (~is_q, is_q):
            off[m] = np.concatenate([[0], np.cumsum(cnt[m])[:-1]]) if m.any() else 0
        sub.trk_key_off = off.astype(np.uint32)
        old_off = fs.trk_key_off[keep_trk].astype(np.int64)

        def take(pool_arrays, mask):
            if not mask.any():
                return [a[:0].copy() for a in pool_arrays]
            idx = np.concatenate([np.arange(o, o + c) for o, c in zip(old_off[mask], cnt[mask])]) if len(set(cnt[mask])) > 1 else \
                (old_off[mask][:, None] + np.arange(int(cnt[mask][0]))[None, :]).reshape(-1)
            return [a[idx] for a in pool_arrays]
        sub.v3_t, sub.v3_val, sub.v3_tension, sub.v3_bias = take([fs.v3_t, fs.v3_val, fs.v3_tension, fs.v3_bias], ~is_q)
        sub.q_t, sub.q_val = take([fs.q_t, fs.q_val], is_q)
        for anims, ts, te in fs.scene_anims:
            a = np.asarray(anims)
            sub.scene_anims.append((new_anim[a[keep_anim[a]]].astype(np.int32), ts, te))
        sub.plays = list(fs.plays)
    return sub
