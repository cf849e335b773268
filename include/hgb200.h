/* hgb200 — C ABI of the B200-native scene-update path for HARFANG scenes.
 *
 * The reference (harfang3d/harfang3d v3.2.7) has no FFI around this path: it is the C++ members of hg::Scene plus the free
 * function CullModelDisplayLists. A maintainer binds this library UNDER that C++ API: hg::Scene keeps its
 * containers and handles; the per-frame functions below replace the bodies of
 *
 *   Scene::StorePreviousWorldMatrices / ReadyWorldMatrices / ComputeWorldMatrices / FixupPreviousWorldMatrices
 *                                                    harfang/engine/scene.cpp:280-304 (driver: Scene::Update :307-315)
 *   Scene::UpdatePlayingAnims -> EvaluateBoundAnim    harfang/engine/scene.cpp:2380-2431, :2154-2181 (transform targets)
 *   Scene::GetModelDisplayLists (enable-flag logic)   harfang/engine/scene.cpp:318-404
 *   CullModelDisplayLists                             harfang/engine/render_pipeline.cpp:2098-2104
 *   bone-matrix loop of _DrawSkinnedModelDisplayLists harfang/engine/render_pipeline.cpp:2172-2181
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on error; the message is in hgb_last_error() (thread-local). Nothing throws.
 *     (The reference's own convention is warn()+neutral value, harfang/engine/scene.cpp:750-766; the C++ facade maps codes to that.)
 *   - one context = one hg::Scene, or one batch of scene instances laid out in one slot space. A context is not thread-safe;
 *     contexts are independent. All work is enqueued on the context's CUDA stream; getters synchronise that stream.
 *   - "slot" = transform component index (the index of Scene::GetTransformWorldMatrices(), scene.cpp:361). Matrices are
 *     hg::Mat4 bit layout: float[12], row-major 3x4 (harfang/foundation/matrix4.h:25). TRS = hg::TransformTRS: pos, rot (Euler
 *     radians, order YXZ), scale = float[9] (harfang/engine/node.h:36-38). Frustum = 6 planes x {x,y,z,w}, order
 *     Top,Bottom,Left,Right,Near,Far = hg::Frustum (harfang/foundation/frustum.h:14-16).
 *   - there is NO CPU fallback: without a CUDA device hgb_create fails.
 */
#ifndef HGB200_H
#define HGB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hgb_ctx hgb_ctx;

#define HGB_ROOT 0xffffffffu      /* parent_slot value of a root transform */
#define HGB_MAX_VIEWS 8           /* frusta tested per entry in one pass (view + 4 PSSM splits + spot = 6 in the reference) */
#define HGB_ENTRY_DISABLED 0x01   /* node has NF_Disabled|NF_InstanceDisabled (harfang/engine/scene.h:166-170) */
#define HGB_KIND_OPAQUE 0
#define HGB_KIND_TRANSPARENT 1
#define HGB_NOT_PLAYING INT64_MIN /* play_t value of a play slot with nothing playing */

typedef struct hgb_options {
	int use_graph; /* 1: capture the per-frame launch sequence in a CUDA graph (default), 0: plain stream launches */
	int row_store; /* profiling A/B switch. 0 (default): a warp's 32 world matrices leave the SM as 3 contiguous 512-byte stores
	                  (shared-memory transpose); 1: every thread stores its own 48-byte row */
	int lookahead; /* profiling A/B switch. L2 software-prefetch distance of the level kernel in positions: 0 = default (one wave of
	                  resident CTAs), -1 = off, n = n positions */
	int tma_level;     /* profiling A/B switch. 0 (default): level kernel with one thread per node, tile rows read with 128-byte LDGs (+ the
	                      L2 look-ahead above); 1: persistent level kernel, node tiles staged with TMA bulk copies, world rows leave as
	                      bulk stores (measured 3 % slower on the 16M-node batch, profiles/README.md r01c) */
	int async_post;    /* the tail of the frame — compaction of the visible lists and the peer gather — runs as its own graph on a second stream
	                      and overlaps the NEXT frame's track sampling and level kernels (visibility bytes are double-buffered by world-buffer
	                      parity; frame N + 2 waits for the tail of frame N, every other entry point joins the pending tails first), so
	                      results and API semantics are unchanged. 0 = default: chosen per frame — on, unless the caller reads the frame's
	                      result on the host (hgb_get_visible*, hgb_peer_result, ...) between every two updates, which leaves the tail
	                      nothing to overlap with: after two such frames the one-graph frame is used (it ends sooner), until an update
	                      finds the previous result unread (with the peer gather only for batches of >= 4 M nodes: on short frames
	                      the tail stays on its own stream). 1 = always on, -1 = off (one graph per frame) */
	int top_walk;      /* the leading levels of the hierarchy go out as ONE launch (k_update_top: every node recomputes its ancestor chain, no
	                      dependency between threads). 0 = default budget (levels are added while sum(nodes x depth) <= 500 K: on the 16M-node batch the same cut as the 250 K measured best there, one launch fewer on c5 / c3, profiles/README.md r02), -1 = off (one
	                      launch per level), n = budget in node evaluations */
	int overlap_sampler; /* 1: tracks are sampled in chunks by target level on a second stream, overlapped with the level kernels. Default 0:
	                        both kernels lean on HBM, running them together measured 1-8 % slower than back to back (profiles r01c) */
	int slice_nodes;   /* with overlap_sampler, levels larger than this are cut into slices of this many nodes, each waiting only for its
	                      own track chunk (0 = no slicing; rounded up to a multiple of 256) */
	int append_slack;  /* > 0: every depth sort (hgb_set_topology) reserves this many free positions behind the last level — the APPEND REGION —
	                      and this many free entries behind each display-entry table, so that hgb_append_nodes / hgb_remove_nodes /
	                      hgb_append_display_entries edit a live scene in place: no re-sort, no table rebuild, no graph capture. The region is
	                      computed by one extra launch per frame (k_update_top over its positions). 0 (default): those calls rebuild */
} hgb_options;
/* Environment switches read by hgb_create, for A/B measurements only (defaults are the fast settings; results are identical in every
 * mode): HGB_PDL=0 launches the frame's kernels with plain stream dependencies instead of programmatic dependent launches;
 * HGB_PDL_LATE=0 puts the level kernel's dependency wait ahead of the sampled values; HGB_PDL_FIRST=0 keeps the first level a plain
 * launch behind the track sampler ("Programmatic dependent launch"). */

/* one bound animation track (BoundToNodeAnim entry, harfang/engine/scene.h:74-88) */
typedef struct hgb_track {
	uint32_t target_slot; /* transform slot of the bound node */
	uint32_t key_off;     /* first key in the pool of its kind */
	uint32_t key_count;
	uint16_t play;        /* play slot whose clock drives it (ScenePlayAnim, scene.h:884-895) */
	uint8_t kind;         /* 0 "Position", 1 "Rotation" Euler vec3, 2 "Scale" (Hermite vec3) ; 3 "Rotation" quaternion (linear) */
	uint8_t pad;
} hgb_track;

const char *hgb_last_error(void);
int hgb_device_count(void);

hgb_ctx *hgb_create(int device, const hgb_options *opt /* nullable */);
void hgb_destroy(hgb_ctx *);
int hgb_set_stream(hgb_ctx *, void *cuda_stream); /* cudaStream_t; NULL = the context's own stream */
int hgb_sync(hgb_ctx *);

/* ---- scene description (edit time) ---- */
/* Transform hierarchy. parent_slot[i] = slot of the parent's transform (node->transform hop already resolved,
 * scene.cpp:268) or HGB_ROOT; used[i]==0 marks holes of the slot list (vector_list.h). Rebuilds the depth-sorted layout. */
int hgb_set_topology(hgb_ctx *, uint32_t n_slots, const uint32_t *parent_slot, const uint8_t *used /* nullable */);
/* Incremental form for SetTransformParent (harfang/engine/component.cpp:184-203) on a few transforms: new parent slot (or HGB_ROOT) per listed slot.
 * When every moved transform stays in its hierarchy level (new parent one level above; roots stay roots) only the parent rows change on the
 * device — no re-sort, no table rebuild, no graph re-capture; *out_rebuilt = 0. Otherwise the layout is rebuilt as by hgb_set_topology and
 * *out_rebuilt = 1 (the caller then re-derives what it derives after hgb_set_topology). */
int hgb_set_parents(hgb_ctx *, uint32_t n, const uint32_t *slot, const uint32_t *parent_slot, int *out_rebuilt /* nullable */);
/* In-place edits of a live scene (spawn / despawn per episode). With hgb_options.append_slack > 0 these cost a few bytes and one tiny
 * launch: no depth sort, no table rebuild, no graph capture; *out_rebuilt = 0. Without slack, or when the fast path does not apply (region
 * full, chain inside the append region longer than 11, node with children removed, second entry of a node, per-instance views), they fall
 * back to the full rebuild and *out_rebuilt = 1. Results are the same either way.
 *   hgb_append_nodes: new LEAF transforms in free slots (Scene::CreateNode + CreateTransform + SetParent, scene.cpp:92-137,
 *     component.cpp:184-193): slot, parent slot (or HGB_ROOT) and TRS (9 floats) per node; a slot may be one past the current end.
 *   hgb_remove_nodes: DestroyNode / DestroyTransform + GarbageCollect (scene.cpp:139-237). Display entries of the node are disabled
 *     for good; children of a removed node become roots (scene.cpp:268-272), which is a rebuild.
 *     Bound tracks are the caller's: a track that targets a removed slot keeps sampling into nothing, but it must be dropped
 *     (hgb_set_tracks) before that slot is handed out again, or it will drive the new node at the next re-bind — the reference's
 *     BoundToNodeAnim holds a generational NodeRef that dies with the node (the facade re-binds on DestroyNode).
 *   hgb_append_display_entries: entries behind the last one of a table, same arguments as hgb_set_display_entries. */
int hgb_append_nodes(hgb_ctx *, uint32_t n, const uint32_t *slot, const uint32_t *parent_slot /* nullable: roots */, const float *trs9, int *out_rebuilt /* nullable */);
int hgb_remove_nodes(hgb_ctx *, uint32_t n, const uint32_t *slot, int *out_rebuilt /* nullable */);
int hgb_append_display_entries(hgb_ctx *, int kind, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, const uint8_t *flags /* nullable */,
	const uint16_t *view /* nullable */, int *out_rebuilt /* nullable */);
int hgb_slot_capacity(hgb_ctx *, uint32_t *out_capacity); /* slots the device arrays hold (topology + append slack) */
/* TRS of slots [first, first+count) — dirty-range upload. */
int hgb_upload_trs(hgb_ctx *, uint32_t first, uint32_t count, const float *trs9);
/* SetNodeWorldMatrix overrides (scene.cpp:768-786) for the NEXT hgb_update only. */
int hgb_set_preset_world(hgb_ctx *, uint32_t n, const uint32_t *slot, const float *mat12);

/* Scene::SetTransformWorldMatrix (is_local = 0, harfang/engine/component.cpp:265-281) or SetTransformLocalMatrix (is_local = 1, :236-248) for
 * a batch of transforms, applied immediately with the reference's sequential semantics: the current world row is written, the local matrix
 * is InverseFast(world[parent]) * world (resp. world = world[parent] * local), and Decompose (foundation/matrix4.cpp:208-252) refreshes
 * pos / Euler rot / scale of the transform (read back with hgb_get_trs). The physics-sync and script use case of SURVEY.md §8f-2. */
int hgb_set_matrices(hgb_ctx *, uint32_t n, const uint32_t *slot, const float *mat12, int is_local);

/* MinMax table: Model::bounds of every (model, list) in use (render_pipeline.h:419-424). */
int hgb_set_bounds(hgb_ctx *, uint32_t n_bounds, const float *minmax6);
/* Display entries of one kind, in the reference's emission order (node index x list index, scene.cpp:330-364).
 * flags: HGB_ENTRY_DISABLED. view: per-entry instance id for hgb_set_instance_views (nullable). */
int hgb_set_display_entries(hgb_ctx *, int kind, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, const uint8_t *flags,
	const uint16_t *view /* nullable */);
int hgb_set_entry_flags(hgb_ctx *, int kind, uint32_t first, uint32_t count, const uint8_t *flags);

/* Skins: one record per (skinned display list, bone), in list order (SkinnedModelDisplayList::mtx_idxs/bones_idxs,
 * render_pipeline.h:596-603, resolved as in scene.cpp:365-394); bind_idx indexes the bind-pose pool. */
int hgb_set_skins(hgb_ctx *, uint32_t n_bones, const uint32_t *mtx_slot, const uint32_t *bind_idx, uint32_t n_bind, const float *bind_pose12);

/* Animation keys and bound tracks. vec3 pool = AnimKeyHermiteT<Vec3> (animation.h:42-46), quat pool = AnimKeyT<Quaternion>. */
int hgb_set_anim_keys(hgb_ctx *, uint32_t n_vec3_keys, const int64_t *v3_t, const float *v3_val3, const float *v3_tension, const float *v3_bias,
	uint32_t n_quat_keys, const int64_t *q_t, const float *q_val4);
int hgb_set_tracks(hgb_ctx *, uint32_t n_tracks, const hgb_track *tracks, uint32_t n_play_slots);

/* "Enable" bool tracks (BoundToNodeAnim::bool_track[NBAT_Enable], harfang/engine/scene.h:74-88), in the reference's evaluation order.
 * Evaluated every hgb_update with EvaluateStep (engine/animation.h:130-139) at the track's play clock: the node of target_slot and all
 * its display entries take HGB_ENTRY_DISABLED = !value before the frame's visibility is computed (Scene::EvaluateBoundAnim
 * scene.cpp:2158-2162 -> EnableNode / DisableNode). kind is ignored; key_off / key_count index the bool key pool given here; play slots
 * are the ones declared by hgb_set_tracks (call it first). The state persists when the track stops, like NF_Disabled. */
int hgb_set_enable_tracks(hgb_ctx *, uint32_t n_tracks, const hgb_track *tracks, uint32_t n_keys, const int64_t *key_t, const uint8_t *key_v);
/* EnableNode / DisableNode from the host on enable-tracked nodes (whose state otherwise lives on the device): sets the node and its entries */
int hgb_set_enable_state(hgb_ctx *, uint32_t n, const uint32_t *slot, const uint8_t *disabled);
/* current state of every enable-tracked node: target slots (ascending) and 1 = disabled */
int hgb_get_enable_state(hgb_ctx *, uint32_t capacity, uint32_t *out_slot, uint8_t *out_disabled, uint32_t *out_count);

/* Views. Either up to HGB_MAX_VIEWS frusta tested against every entry, or one frustum per scene instance (entry -> view[]). */
int hgb_set_views(hgb_ctx *, uint32_t n_views, const float *frustum24);
int hgb_set_instance_views(hgb_ctx *, uint32_t n_instances, const float *frustum24, const uint32_t *inst_entry_start /* n_instances+1, kind 0 */);

/* ---- per frame ---- */
/* Scene::Update body: swap current/previous worlds, clear updated flags, apply presets, sample the tracks at play_t[]
 * (clocks are stepped by the caller exactly as scene.cpp:2386-2413 — int64, host side), propagate world matrices level by level.
 * fused_cull != 0: the visibility of every display entry is evaluated in the same pass and compacted (hgb_cull is implied). */
int hgb_update(hgb_ctx *, uint32_t n_play, const int64_t *play_t, int fused_cull);
/* CullModelDisplayLists against the current views using the stored world matrices. */
int hgb_cull(hgb_ctx *);
/* CullModelDisplayLists (render_pipeline.cpp:2098-2104) for an arbitrary caller-owned list, exactly like the reference's free function:
 * entry e uses the current world matrix of transform slot mtx_slot[e] and the MinMax bounds6[bounds_idx[e]]; the surviving entry
 * indices are written to out_idx in order (capacity n) and their number to *out_count. Independent of the entry tables above. */
int hgb_cull_list(hgb_ctx *, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, uint32_t n_bounds, const float *bounds6,
	const float *frustum24, uint32_t *out_idx, uint32_t *out_count);
/* OPT-IN culling of skinned display lists. The reference never culls them (harfang/engine/scene_forward_pipeline.cpp:560-562 "FIXME cull skinned
 * models !!!") and neither does anything here unless this is called. A skinned list (<= 32 bones, render_pipeline.h:38) survives unless the
 * union over its bones of bounds6[bounds_idx[i]] carried through world[mtx_slot] * bind_pose[bind_idx] (the skinning matrix of
 * render_pipeline.cpp:2172-2173) is V_Outside — a conservative bound of every skinned vertex. List i owns the bone records
 * [bone_off[i], bone_off[i+1]); surviving list indices come back in order. Stateless, like hgb_cull_list. */
int hgb_cull_skinned_list(hgb_ctx *, uint32_t n_lists, const uint32_t *bone_off /* n_lists + 1 */, const uint32_t *mtx_slot, const uint32_t *bind_idx,
	uint32_t n_bind, const float *bind_pose12, const uint32_t *bounds_idx /* per list */, uint32_t n_bounds, const float *bounds6, const float *frustum24,
	uint32_t *out_idx, uint32_t *out_count);
/* ComputeModelDisplayListSortKeys (harfang/engine/scene_forward_pipeline.cpp:438-483, ComputeSortKey render_pipeline.cpp:89-94) for the
 * survivors of list (kind, view) of the last cull: one 1-mm fixed-point key per visible entry, from the closest view-space z of the 8
 * corners of its MinMax. view_mat12 = ViewState::view (hg::Mat4). Keys are returned in visible-list order by hgb_get_sort_keys. */
int hgb_sort_keys(hgb_ctx *, int kind, uint32_t view, const float *view_mat12);
/* Scene::GetMinMax (harfang/engine/scene.cpp:1982) / GetNodesMinMax (engine/node.cpp:265-285) as a device reduction over the current world
 * matrices: per segment, the union over its objects of world[mtx_slot] * object_bounds (Node::GetMinMax node.cpp:106-119,
 * Scene::GetObjectMinMax component.cpp:778-797). hgb_set_objects describes the object-carrying nodes once (edit time): mtx_slot[i] is the
 * node's transform slot or HGB_ROOT when it has none (identity world, like Transform::GetWorld on an invalid ref), obounds_idx[i] indexes
 * obounds6 = one MinMax per model (the union of its lists' bounds; nodes whose model is invalid or has no bounds are left out by the caller,
 * as the reference skips them), seg[i] (nullable) is the scene instance of the object, n_seg the number of segments.
 * hgb_minmax enqueues the reduction; hgb_get_minmax returns per segment {mn.xyz, mx.xyz} and the number of contributing nodes
 * (0 = GetNodesMinMax returns false, the 6 floats are then unspecified). */
int hgb_set_objects(hgb_ctx *, uint32_t n_obj, const uint32_t *mtx_slot, const uint32_t *obounds_idx, uint32_t n_obounds, const float *obounds6,
	const uint32_t *seg /* nullable */, uint32_t n_seg);
int hgb_minmax(hgb_ctx *);
int hgb_get_minmax(hgb_ctx *, float *out6 /* n_seg x 6 */, uint32_t *out_count /* n_seg, nullable */);
/* Bone matrices of every skin record (and the previous-frame twin). */
int hgb_bone_matrices(hgb_ctx *, int with_previous);

/* ---- results (device -> host; each call synchronises the stream) ---- */
int hgb_get_worlds(hgb_ctx *, uint32_t first, uint32_t count, float *out12, int previous);
int hgb_get_trs(hgb_ctx *, uint32_t first, uint32_t count, float *out9);
int hgb_get_visible_count(hgb_ctx *, int kind, uint32_t view, uint32_t *out_count);
/* surviving entry indices (into the table given to hgb_set_display_entries), ascending = reference order */
int hgb_get_visible(hgb_ctx *, int kind, uint32_t view, uint32_t *out_idx, uint32_t capacity, uint32_t *out_count);
/* the same list in a pinned buffer OWNED by the context (valid until the next call for this kind, or hgb_destroy): count and list come down
 * in one round trip and land by plain DMA whatever memory the caller would have offered */
int hgb_get_visible_ptr(hgb_ctx *, int kind, uint32_t view, const uint32_t **out_idx, uint32_t *out_count);
/* Optional: one opaque 16-byte record per display entry of a kind (what the caller wants back per survivor — the facade stores its
 * hg::ModelDisplayList, harfang/engine/render_pipeline.h:577-582), and the visible list of (kind, view) returned AS those records, in order:
 * gathered on the device, one round trip, pinned buffer owned by the context (valid until the next call for this kind). */
int hgb_set_entry_records(hgb_ctx *, int kind, uint32_t first, uint32_t count, const void *rec16); /* contiguous from 0: the whole table, or the tail behind the records already set */
int hgb_get_visible_records_ptr(hgb_ctx *, int kind, uint32_t view, const void **out_rec16, uint32_t *out_count);
/* ... pipelined: returns as soon as the count is known; the records arrive in chunks of `chunk_records` (>= 1024), chunk i has landed in the
 * pinned buffer when hgb_visible_records_wait(ctx, kind, i) returns 0 (callable from any thread), so a caller copies chunk i into its own
 * container while chunk i + 1 is still on the link (the facade's GetVisibleModelDisplayLists does). */
int hgb_get_visible_records_begin(hgb_ctx *, int kind, uint32_t view, uint32_t chunk_records, const void **out_rec16, uint32_t *out_count);
int hgb_visible_records_wait(hgb_ctx *, int kind, uint32_t chunk);
/* instance mode: exclusive offsets of every instance's survivors inside the kind-0 list, n_instances+1 values */
int hgb_get_instance_offsets(hgb_ctx *, uint32_t *out_offsets);
int hgb_get_bone_matrices(hgb_ctx *, float *out16, int previous);
int hgb_get_sort_keys(hgb_ctx *, int kind, uint32_t view, uint32_t *out_keys, uint32_t capacity, uint32_t *out_count);

/* ---- multi-GPU: gather of the visible lists to one rank over NVLink / NVSwitch peer memory ----
 * One context per GPU (one process per GPU); scene instances are sharded, the only exchange is this gather of the RESULT. After
 * hgb_peer_connect every fused hgb_update also: stores the survivors of list (kind 0, view 0), globalised with the rank's first entry
 * index, straight from the compaction kernel into this rank's region of the destination GPU's inbox (peer stores), publishes count and a
 * step flag, and — on the destination — waits for the step's flags, concatenates the regions in rank order and acknowledges. All of it
 * is enqueued on the contexts' streams: no host synchronisation, no host-side sizes. Every rank must call hgb_update the same number of
 * times. Setup: every rank calls hgb_peer_init (entries_per_rank = display entries of kind 0 on every rank = region capacities;
 * aux_words_per_rank = scene instances per rank in instance mode, else NULL), the handles are exchanged by the caller (64 opaque bytes
 * each, e.g. torch.distributed.all_gather_object), then hgb_peer_connect(inbox handle of dst, mailbox handles of all ranks). */
typedef struct hgb_peer_handle {
	unsigned char bytes[64]; /* cudaIpcMemHandle_t */
} hgb_peer_handle;
int hgb_peer_init(hgb_ctx *, uint32_t rank, uint32_t world, uint32_t dst, const uint32_t *entries_per_rank, const uint32_t *aux_words_per_rank /* nullable */,
	hgb_peer_handle *inbox_out /* meaningful on dst */, hgb_peer_handle *mailbox_out);
int hgb_peer_connect(hgb_ctx *, const hgb_peer_handle *inbox_of_dst, const hgb_peer_handle *mailboxes /* world handles, read on dst */);
/* undo hgb_peer_init / hgb_peer_connect (also after a setup that failed on this or another rank): waits for the frames in flight, unmaps
 * and frees the peer buffers; later frames run without the gather */
int hgb_peer_close(hgb_ctx *);
/* frames counted so far and the error word (0 = ok; 1 = a sender timed out waiting for an acknowledgement, 2 + r = the destination timed
 * out waiting for rank r) */
int hgb_peer_status(hgb_ctx *, uint32_t *out_step, uint32_t *out_error);
/* destination only, after a frame: rank offsets (world + 1), the concatenated global entry indices, and the aux words (per-instance
 * exclusive offsets of the whole batch followed by the total, like hgb_get_instance_offsets of one context holding every instance) */
int hgb_peer_result(hgb_ctx *, uint32_t *out_rank_off, uint32_t *out_idx /* nullable */, uint32_t idx_capacity, uint32_t *out_aux /* nullable */,
	uint32_t aux_capacity);

/* ---- introspection for benchmarks / interop ---- */
/* device pointers: 0 worlds, 1 previous worlds, 2 visible list (kind 0, view 0), 3 visible count (kind 0, view 0), 4 bone matrices,
 * 5 instance offsets, 6 TRS SoA, 7 gathered visible list (peer gather destination), 8 its rank offsets */
void *hgb_device_ptr(hgb_ctx *, int which);
/* ---- allocation-free entry point ----
 * No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it). Order-preserving compaction of one view's bit of a
 * visibility-byte array — what k_compact does behind hgb_update / hgb_cull — on the CALLER's device buffers: d_vis[n] (one byte per
 * entry, bit v set = survives view v; 16-byte aligned, no padding needed), d_out[out_capacity] receives the surviving entry indices in
 * ascending order (writes stop at out_capacity), *d_count the total number of survivors (may exceed out_capacity). Enqueued on
 * `cuda_stream` (cudaStream_t, NULL = default stream) of `device`, which it makes current; returns once enqueued — synchronise the
 * stream before reading. Calls no cudaMalloc / cudaFree / allocating routine: the kernel's scratch (tile ticket + look-back status words)
 * is a static __device__ array of 8.25 KB that every GPU loading the library holds until exit, which bounds n at
 * HGB_NOALLOC_MAX_ENTRIES (larger n is refused with an error). All calls on a device SHARE that array: each call clears what it uses on
 * its stream, so back-to-back calls on one stream are fine, but calls whose GPU work overlaps (different streams or threads) race. */
#define HGB_NOALLOC_MAX_ENTRIES (1u << 24)
int hgb_compact_visible_noalloc(int device, const uint8_t *d_vis, uint32_t n, uint32_t view_bit, uint32_t *d_out, uint32_t out_capacity, uint32_t *d_count,
	void *cuda_stream);

int hgb_frame_was_async(hgb_ctx *);       /* 1: the last hgb_update ran with the asynchronous tail (hgb_options.async_post), 0: as one graph */
uint64_t hgb_rebuild_count(hgb_ctx *);  /* depth sorts + entry / track table rebuilds + graph captures so far: an in-place edit leaves it unchanged */
uint64_t hgb_launch_count(hgb_ctx *);   /* kernels launched by this context so far (graph launches count their kernel nodes) */
uint32_t hgb_level_count(hgb_ctx *);
int hgb_is_identity_order(hgb_ctx *);   /* 1 when slot order already is the depth-sorted order (no shuffle kernels needed) */
/* CUDA-event stopwatch on the context's stream: begin records an event, end records one, synchronises and returns the ms between. */
int hgb_time_begin(hgb_ctx *);
int hgb_time_end(hgb_ctx *, float *out_ms);
/* Per-kernel CUDA-event timing. While on, frames are enqueued as plain stream launches (no graph) and every kernel launch is
 * bracketed by an event pair on the context's stream. hgb_get_profile synchronises and returns the accumulated milliseconds and
 * launch counts per kernel class (arrays of HGB_K_COUNT). */
#define HGB_K_SAMPLE_TRACKS 0
#define HGB_K_UPDATE_LEVEL 1
#define HGB_K_COMPACT 2
#define HGB_K_SEGMENT_RANKS 3
#define HGB_K_CULL_FLAGS 4
#define HGB_K_BONE_MATRICES 5
#define HGB_K_SORT_KEYS 6
#define HGB_K_MINMAX 7
#define HGB_K_COUNT 8
int hgb_set_profiling(hgb_ctx *, int on);
int hgb_get_profile(hgb_ctx *, double *ms_by_class, uint64_t *launches_by_class, int reset);
/* Write `bytes` bytes of a scratch buffer (allocated on first use) — L2 flush between timed iterations. */
int hgb_flush_l2(hgb_ctx *, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif
