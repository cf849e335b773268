// sm_100a kernels of the scene-update path. All of them are HBM-bound streaming / gather / compaction kernels (no tensor
// cores: ~1.3 flop/B). Data layout: everything on the device is stored in DEPTH-SORTED POSITION order
// ("pos"): position p of a transform slot is its rank in a level-by-level (BFS) ordering with siblings adjacent, so a level
// is a contiguous range, every parent has a smaller position than its children, and sibling threads hit the same parent lines.
#pragma once
#include "hgb_math.cuh"

namespace hgb {

constexpr uint32_t kRoot = 0xffffffffu;
constexpr uint32_t kDead = 0xfffffffeu; // parent row of a free position of the append region (never seen by the level kernels)
constexpr int kMaxViews = 8;

// ---- node tiles -------------------------------------------------------------------------------------------------------------
// Per-node inputs live in TILES of 32 consecutive positions: 12 rows of 32 words = 1536 contiguous bytes, so one warp reads its
// 32 nodes as 12 fully coalesced 128-byte rows at immediate offsets from a single address:
//   rows 0..8  px py pz rx ry rz sx sy sz   (hg::TransformTRS, engine/node.h:36-38)
//   row  9     parent position (kRoot for roots)
//   row  10    first display entry of the node: bounds index (bits 0..23) | kEnt* flags | animated-triple bits
//   row  11    index of that entry in its kind's table
// = exactly the 48 algorithmic bytes a node contributes to the fused pass. Nodes with more than one display entry (multi-list
// models) continue in an overflow CSR (ent_off / pent), read only when kEntMore is set.
constexpr uint32_t kTileRows = 12, kTileWords = kTileRows * 32;
constexpr uint32_t kRowParent = 9, kRowEntMeta = 10, kRowEntIdx = 11;
__host__ __device__ __forceinline__ size_t tile_word(uint32_t p, uint32_t row) { return size_t(p >> 5) * kTileWords + row * 32u + (p & 31u); }

// entry meta word: bounds index in the low 24 bits, flags above
constexpr uint32_t kEntDisabled = 1u << 24;        // node has NF_Disabled|NF_InstanceDisabled
constexpr uint32_t kLaneAnimated = 1u << 25;       // bits 25/26/27: the pos / rot / scale triple of this node is driven by bound tracks:
                                                   // its x row holds an index into anim_val[] (float4) instead of a value (see k_sample_tracks)
constexpr uint32_t kLaneAnimatedMask = 7u << 25;
constexpr uint32_t kEntHas = 1u << 29;             // row 10/11 hold an entry
constexpr uint32_t kEntMore = 1u << 30;            // further entries of this node in the overflow CSR
constexpr uint32_t kEntKindTransparent = 1u << 31; // which of the two entry tables (row 10; bit 31 of PosEntry::x in the CSR)

// View frusta live in device memory (6 float4 per view: Top,Bottom,Left,Right,Near,Far, foundation/frustum.h:14; w holds -2*w) so that a camera
// move is a 96-byte copy and never invalidates the captured frame graph.
struct Views {
	const float4 *planes;   // 6 float4 per view
	int n_views;
	const float4 *planes_x; // the same planes expanded for the packed test (survives_frustum_x): 12 float4 per view
};

struct SceneDev {
	const uint32_t *tiles;      // node tiles (see above)
	const float4 *anim_val;     // sampled pos / rot / scale triples of track-driven nodes (written coalesced by k_sample_tracks)
	float *world;               // AoS 12 floats per position, 16-B aligned (== hg::Mat4 bit layout)
	const uint8_t *preset;      // nullable: 1 -> world[p] was pre-set this frame (Scene::SetNodeWorldMatrix), keep it
	// fused cull
	const uint32_t *ent_off;    // overflow CSR: n_pos + 1
	const uint2 *pent;          //   {entry index | kind bit 31, meta}
	const float *bounds6;       // MinMax table, 6 floats per (model, list)
	uint8_t *vis0, *vis1;       // per entry table (opaque, transparent): bit v set -> survives view v
	const float4 *inst_planes;  // per-instance frusta (6 float4 each) or null
	const float4 *inst_planes_x; // ... expanded (12 float4 each)
	const uint16_t *pos_view;   // per position: which instance frustum (only with inst_planes)
	int transpose_store;        // 1: world rows leave the SM as 3 contiguous 512-byte warp stores (shared-memory transpose)
	uint32_t lookahead;         // positions: a warp asks L2 for the tile (and its parents' rows) this far ahead of its own; 0 = off
	int late_wait;              // where griddepcontrol.wait sits: 0 ahead of the sampled values; 1 behind the local matrix (levels behind a level);
	                            // 2 like 0 and dependents are released only behind it (the level chained to the track sampler)
};

// mbarrier + 1-D bulk copy (TMA) helpers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
	uint32_t ok;
	asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
	return ok != 0;
}
__device__ __forceinline__ void bulk_load(uint32_t dst_smem, const void *src, uint32_t bytes, uint32_t bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar)
				 : "memory");
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// visibility mask of one display entry against the shared views (shared memory) or the node's instance frustum
// `planes_x`: expanded planes (2 float4 per plane) in shared memory
template <bool PER_INSTANCE>
__device__ __forceinline__ uint32_t entry_visibility(const SceneDev &s, const Mat4 &world, uint32_t meta, const float4 *planes_x, int n_views, uint32_t p) {
	if (meta & kEntDisabled)
		return 0;
	const Aabb bb = minmax_transform(world, s.bounds6 + size_t(meta & 0xffffffu) * 6);
	if (PER_INSTANCE)
		return survives_frustum_x(s.inst_planes_x + size_t(s.pos_view[p]) * 12, bb) ? 1u : 0u;
	uint32_t bits = survives_frustum_x(planes_x, bb) ? 1u : 0u;
	for (int v = 1; v < n_views; ++v)
		bits |= survives_frustum_x(planes_x + v * 12, bb) ? (1u << v) : 0u;
	return bits;
}

// fill the CTA's expanded copy of the shared view frusta (call before a __syncthreads)
__device__ __forceinline__ void stage_planes(float4 *s_planes, const Views &views) {
	if (threadIdx.x < uint32_t(views.n_views) * 6u)
		expand_plane(views.planes[threadIdx.x], s_planes[threadIdx.x * 2], s_planes[threadIdx.x * 2 + 1]);
}

// ---- per-node pieces shared by the level kernels ----------------------------------------------------------------------------
// TRS of one node: its nine tile rows, with the track-driven triples replaced by the values sampled this frame (k_sample_tracks).
// The position triple comes back as a pending gather (`pos_slot`), consumed only once the rotation is built: the translation is the
// last thing TransformationMat4 needs, so the L2 round trip of that gather hides behind the three sincos.
struct NodeTrs {
	float px, py, pz, ex, ey, ez, sx, sy, sz;
};
__device__ __forceinline__ NodeTrs load_trs(const SceneDev &s, const uint32_t *t, uint32_t meta) {
	NodeTrs n;
	n.px = __uint_as_float(t[0]), n.py = __uint_as_float(t[32]), n.pz = __uint_as_float(t[64]);
	n.ex = __uint_as_float(t[96]), n.ey = __uint_as_float(t[128]), n.ez = __uint_as_float(t[160]);
	n.sx = __uint_as_float(t[192]), n.sy = __uint_as_float(t[224]), n.sz = __uint_as_float(t[256]);
	if (meta & kLaneAnimatedMask) { // track-driven triples: the x row holds the anim_val slot
		if (meta & (kLaneAnimated << 1)) {
			const float4 v = s.anim_val[__float_as_uint(n.ex)];
			n.ex = v.x, n.ey = v.y, n.ez = v.z;
		}
		if (meta & (kLaneAnimated << 2)) {
			const float4 v = s.anim_val[__float_as_uint(n.sx)];
			n.sx = v.x, n.sy = v.y, n.sz = v.z;
		}
	}
	return n;
}
// local = TransformationMat4(pos, rot, scl) (foundation/matrix4.cpp:176-183), in the row-pair form mat4_mul_p consumes
__device__ __forceinline__ Mat4p node_local(const SceneDev &s, const uint32_t *t, uint32_t meta) {
	const NodeTrs n = load_trs(s, t, meta);
	float4 vp = make_float4(n.px, n.py, n.pz, 0.f);
	if (meta & kLaneAnimated)
		vp = s.anim_val[__float_as_uint(n.px)];
	return transformation_mat4p(vp.x, vp.y, vp.z, n.ex, n.ey, n.ez, n.sx, n.sy, n.sz);
}

// world rows of a warp's 32 consecutive positions: through a shared-memory transpose they leave the SM as 3 contiguous 512-byte stores
__device__ __forceinline__ void store_world_rows(const SceneDev &s, float4 *so, const Mat4 &world, bool live, uint32_t p, uint32_t lane, uint32_t end) {
	if (s.transpose_store) {
		if (live) {
			so[lane * 3 + 0] = make_float4(world.m[0][0], world.m[0][1], world.m[0][2], world.m[0][3]);
			so[lane * 3 + 1] = make_float4(world.m[1][0], world.m[1][1], world.m[1][2], world.m[1][3]);
			so[lane * 3 + 2] = make_float4(world.m[2][0], world.m[2][1], world.m[2][2], world.m[2][3]);
		}
		__syncwarp();
		const uint32_t p0 = p - lane;                              // first position of the warp
		const uint32_t n_f4 = (min(end, p0 + 32u) - p0) * 3u;       // float4s this warp owns
		float4 *dst = reinterpret_cast<float4 *>(s.world) + size_t(p0) * 3;
#pragma unroll
		for (int j = 0; j < 3; ++j)
			if (lane + 32u * j < n_f4)
				dst[lane + 32u * j] = so[lane + 32u * j];
	} else if (live) {
		store_mat4(s.world + size_t(p) * 12, world);
	}
}

// GetModelDisplayLists flag logic + CullModelDisplayLists for every display entry attached to position p, world matrix in registers
// further display entries of a node (multi-list models): rare, kept out of line — and fed with values, not with a reference to the
// kernel's parameter block, which would force the whole block into local memory
struct MoreEntries {
	const uint32_t *ent_off;
	const uint2 *pent;
	const float *bounds6;
	uint8_t *vis0, *vis1;
	const float4 *inst_planes_x;
	const uint16_t *pos_view;
};
template <bool PER_INSTANCE>
__device__ __noinline__ void cull_more_entries(MoreEntries m, Mat4 world, const float4 *s_planes, int n_views, uint32_t p) {
	SceneDev s{};
	s.bounds6 = m.bounds6, s.inst_planes_x = m.inst_planes_x, s.pos_view = m.pos_view;
	const uint32_t e0 = m.ent_off[p], e1 = m.ent_off[p + 1];
	for (uint32_t k = e0; k < e1; ++k) {
		const uint2 rec = m.pent[k];
		const uint32_t b = entry_visibility<PER_INSTANCE>(s, world, rec.y, s_planes, n_views, p);
		((rec.x >> 31) ? m.vis1 : m.vis0)[rec.x & 0x7fffffffu] = uint8_t(b);
	}
}
template <bool PER_INSTANCE>
__device__ __forceinline__ void cull_node_entries(const SceneDev &s, const Mat4 &world, uint32_t meta, uint32_t eidx, const float4 *s_planes, int n_views, uint32_t p) {
	const uint32_t bits = entry_visibility<PER_INSTANCE>(s, world, meta, s_planes, n_views, p);
	((meta & kEntKindTransparent) ? s.vis1 : s.vis0)[eidx] = uint8_t(bits);
	if (meta & kEntMore)
		cull_more_entries<PER_INSTANCE>(MoreEntries{s.ent_off, s.pent, s.bounds6, s.vis0, s.vis1, s.inst_planes_x, s.pos_view}, world, s_planes, n_views, p);
}

// ---- world-matrix propagation, one level per launch ---------------------------------------------------------------------
// Replaces Scene::ComputeTransformWorldMatrix (engine/scene.cpp:263-277) for the positions [begin, end) of one hierarchy
// level: local = TransformationMat4(pos, rot, scl); world = parent ? world[parent] * local : local. Parents were written
// by the previous launch. With FUSED the freshly computed world matrix is tested, still in registers, against the
// view frusta for every display entry attached to the position (GetModelDisplayLists flag logic + CullModelDisplayLists,
// engine/scene.cpp:330-364, engine/render_pipeline.cpp:2098-2104) and only a 1-byte visibility mask per entry is written.
// `begin` is a multiple of 32 relative to the grid (blockDim = 256), so a warp covers 32 consecutive positions.
// MINB = resident CTAs per SM the register allocation is held to (A/B: 6 -> 40 registers, 5 -> 48).
template <bool FUSED, bool PER_INSTANCE, int MINB>
__global__ void __launch_bounds__(256, MINB) k_update_level(SceneDev s, Views views, uint32_t begin, uint32_t end) {
	__shared__ float4 s_out[8][96]; // per warp: 32 world matrices = 1536 B, for the transposed store
	__shared__ __align__(128) float4 s_planes[kMaxViews * 12];
	// The shared view frusta (host-expanded form) are needed only by the cull tail: one thread starts a bulk copy (TMA, completion counted
	// on an mbarrier: no registers, nobody waits) and every warp merely tests the barrier's phase when it reaches the tail, hundreds of
	// instructions later. The only CTA-wide barrier is the one right here, which publishes the initialised mbarrier while all warps are
	// still in step; a __syncthreads() ahead of the tail instead cost 5.7 % of the stall samples (warps of a CTA drift apart).
	__shared__ __align__(8) uint64_t s_bar;
	if (FUSED && !PER_INSTANCE) {
		if (threadIdx.x == 0) {
			const uint32_t bar = smem_u32(&s_bar), bytes = uint32_t(views.n_views) * 192u;
			mbar_init(bar, 1);
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
			mbar_arrive_expect_tx(bar, bytes); // (0 views: the phase completes right here)
			if (bytes)
				bulk_load(smem_u32(s_planes), views.planes_x, bytes, bar);
		}
		__syncthreads();
	}
	// Programmatic dependent launch: the next level's grid may become resident while this one drains — ahead of its own
	// griddepcontrol.wait it reads only edit-time rows and sampled values — which hides the launch latency, the first DRAM round trip and
	// the local-matrix arithmetic of the dependent chain of levels. The level chained to the track sampler (late_wait == 2) releases its
	// dependents only behind its wait, so that no later level can run ahead of the sampler.
	if (s.late_wait != 2)
		asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t p = begin + blockIdx.x * blockDim.x + threadIdx.x;
	const bool live = p < end;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	if (__ballot_sync(0xffffffffu, live) == 0)
		return;
	// Software prefetch into L2: CTAs are dispatched in position order, so the tile `lookahead` positions ahead belongs to a CTA that
	// starts about one CTA lifetime from now. Its 12 rows are requested by 12 lanes here, which turns the first (and otherwise fully
	// exposed) DRAM latency of that CTA into an L2 hit.
	if (s.lookahead != 0 && lane < kTileRows && p - lane + s.lookahead < end)
		prefetch_l2(s.tiles + tile_word((p - lane + s.lookahead) & ~31u, lane));

	Mat4 world;
	uint32_t meta = 0, eidx = 0;
	if (live) {
		const uint32_t *t = s.tiles + tile_word(p, 0);
		meta = t[kRowEntMeta * 32];
		if (FUSED)
			eidx = t[kRowEntIdx * 32];
		if (s.preset && s.preset[p]) {
			// a preset row (SetNodeWorldMatrix) is kept as it is — but the thread still passes the dependency wait: a grid in which no
			// thread waits could complete ahead of its predecessor and break the transitive completion the chain of levels relies on
			asm volatile("griddepcontrol.wait;" ::: "memory");
			if (s.late_wait == 2)
				asm volatile("griddepcontrol.launch_dependents;");
			world = load_mat4(s.world + size_t(p) * 12);
		} else {
			const uint32_t par = t[kRowParent * 32];
			// everything above reads edit-time data; from here on sampled values and the previous level's world rows are read (no-op
			// without a programmatic dependency)
			if (s.late_wait != 1) {
				asm volatile("griddepcontrol.wait;" ::: "memory");
				if (s.late_wait == 2) // the next level reads sampled values ahead of ITS wait: it may only start once the sampler is complete
					asm volatile("griddepcontrol.launch_dependents;");
			}
			if (par != kRoot) { // start the parent row on its way while the trigonometry runs (no registers held)
				if (s.late_wait == 1) { // not into L1 ahead of the wait: the row may not have been written yet and L1 would keep the old line
					prefetch_l2(s.world + size_t(par) * 12);
					prefetch_l2(s.world + size_t(par) * 12 + 11);
				} else {
					prefetch_l1(s.world + size_t(par) * 12);
					prefetch_l1(s.world + size_t(par) * 12 + 11);
				}
			}
			const Mat4p local = node_local(s, t, meta);
			if (s.late_wait == 1)
				asm volatile("griddepcontrol.wait;" ::: "memory");
			if (par != kRoot)
				world = mat4_mul_p(load_mat4(s.world + size_t(par) * 12), local);
			else
				world = unpack_mat4(local);
		}
	}

	// store the world rows (a preset row is rewritten with its own value)
	store_world_rows(s, s_out[warp], world, live, p, lane, end);

	if (FUSED && !PER_INSTANCE) // the frusta have landed long ago: one successful phase test
		while (!mbar_try_wait(smem_u32(&s_bar), 0u)) {
		}
	if (FUSED && live && (meta & kEntHas))
		cull_node_entries<PER_INSTANCE>(s, world, meta, eidx, s_planes, views.n_views, p);
}

// ---- the top of the hierarchy in ONE launch ---------------------------------------------------------------------------------
// The upper levels of a hierarchy hold 1 ... ~10^5 nodes each; one launch per level costs ~7 µs apiece there (launch ramp plus the
// dependent tile -> parent -> math chain), far more than their bytes. This kernel covers the positions [0, end) of the first few
// levels with NO dependency between its threads: every node walks up its ancestor chain (parent rows are edit-time data) and
// multiplies the local matrices down again, world = ((L_root * L_1) * L_2) ... * L_node — the very sequence of operations
// ComputeTransformWorldMatrix performs through its recursion (scene.cpp:263-277: world[parent] is itself world[grandparent] * local),
// so every row is bit-identical to the level-by-level result; it is only recomputed by each descendant instead of read back. The
// host bounds the redundant work (sum over the covered levels of nodes x depth, hgb_options.top_walk) and the chain length
// (kMaxWalk). A chain stops at a preset ancestor (SetNodeWorldMatrix), whose row k_apply_presets wrote ahead of the frame.
//
// The same kernel computes the APPEND REGION behind the last level (hgb_append_nodes): positions [begin, end) hold nodes added since
// the last depth sort, in no particular order, and free positions (parent row kDead). It runs after the last level, so a chain stops at
// the first ancestor below `base` — a sorted position, final for this frame — and only ancestors inside the region are recomputed.
constexpr int kMaxWalk = 12;

template <bool FUSED, bool PER_INSTANCE>
__global__ void __launch_bounds__(256, 6) k_update_top(SceneDev s, Views views, uint32_t begin, uint32_t end, uint32_t base) {
	__shared__ float4 s_out[8][96];
	__shared__ float4 s_planes[kMaxViews * 12];
	if (FUSED && !PER_INSTANCE) {
		if (threadIdx.x < 32u)
			for (uint32_t i = threadIdx.x; i < uint32_t(views.n_views) * 12u; i += 32u)
				s_planes[i] = views.planes_x[i];
		__syncthreads();
	}
	// chained behind the track sampler (or plain): it reads sampled values throughout, so it waits first thing and releases its
	// dependents only behind the wait (they read sampled values ahead of THEIR wait)
	asm volatile("griddepcontrol.wait;" ::: "memory");
	asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t p = begin + blockIdx.x * blockDim.x + threadIdx.x;
	const bool live = p < end;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	if (__ballot_sync(0xffffffffu, live) == 0)
		return;
	Mat4 world;
	uint32_t meta = 0, eidx = 0;
	if (live) {
		const uint32_t *t = s.tiles + tile_word(p, 0);
		uint32_t q = t[kRowParent * 32];
		meta = t[kRowEntMeta * 32];
		if (FUSED)
			eidx = t[kRowEntIdx * 32];
		if (q == kDead) { // free position of the append region: a zero row, no entry
			meta = 0;
#pragma unroll
			for (int i = 0; i < 3; ++i)
#pragma unroll
				for (int j = 0; j < 4; ++j)
					world.m[i][j] = 0.f;
		} else if (s.preset && s.preset[p]) {
			world = load_mat4(s.world + size_t(p) * 12);
		} else {
			uint32_t chain[kMaxWalk]; // ancestors that must be recomputed, nearest first
			int n = 0;
			while (q != kRoot && q >= base && !(s.preset && s.preset[q]) && n < kMaxWalk) {
				chain[n++] = q;
				q = s.tiles[tile_word(q, kRowParent)];
			}
			bool have = q != kRoot; // stopped at a preset ancestor or below the region: its row is the base
			if (have)
				world = load_mat4(s.world + size_t(q) * 12);
			for (int i = n - 1; i >= 0; --i) {
				const uint32_t a = chain[i];
				const uint32_t *ta = s.tiles + tile_word(a, 0);
				const Mat4p la = node_local(s, ta, ta[kRowEntMeta * 32]);
				world = have ? mat4_mul_p(world, la) : unpack_mat4(la);
				have = true;
			}
			const Mat4p local = node_local(s, t, meta);
			world = have ? mat4_mul_p(world, local) : unpack_mat4(local);
		}
	}
	store_world_rows(s, s_out[warp], world, live, p, lane, end);
	if (FUSED && live && (meta & kEntHas))
		cull_node_entries<PER_INSTANCE>(s, world, meta, eidx, s_planes, views.n_views, p);
}

// ---- stand-alone cull predicate (entry order) -------------------------------------------------------------------------------
// CullModelDisplayLists for a list that is culled again without a scene update (extra views, shadow passes:
// engine/forward_pipeline.cpp:331-332,380-381).
struct CullDev {
	const uint32_t *mtx_pos; // per entry
	const uint32_t *meta;    // per entry: bounds idx | flags
	const uint16_t *view;    // per entry instance id (PER_INSTANCE) or null
	const float *world;
	const float *bounds6;
	const float4 *inst_planes;
	uint8_t *vis;
	uint32_t n;
};

template <bool PER_INSTANCE> __global__ void __launch_bounds__(256) k_cull_flags(CullDev c, Views views) {
	__shared__ float4 s_planes[kMaxViews * 6];
	if (!PER_INSTANCE) {
		if (threadIdx.x < uint32_t(views.n_views) * 6u)
			s_planes[threadIdx.x] = views.planes[threadIdx.x];
		__syncthreads();
	}
	const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= c.n)
		return;
	const uint32_t meta = c.meta[e];
	uint32_t bits = 0;
	if (!(meta & kEntDisabled)) {
		const Mat4 world = load_mat4(c.world + size_t(c.mtx_pos[e]) * 12);
		const Aabb bb = minmax_transform(world, c.bounds6 + size_t(meta & 0xffffffu) * 6);
		if (PER_INSTANCE) {
			bits = survives_frustum(c.inst_planes + size_t(c.view[e]) * 6, bb) ? 1u : 0u;
		} else {
			for (int v = 0; v < views.n_views; ++v)
				bits |= survives_frustum(views.planes + v * 6, bb) ? (1u << v) : 0u;
		}
	}
	c.vis[e] = uint8_t(bits);
}

// ---- order-preserving stream compaction --------------------------------------------------------------------------------------
// The reference's cull is std::remove_if + resize (stable). Single pass: each CTA takes the next tile (ticket counter, so
// tiles are started in order), popcounts its 8192 visibility bits, publishes its aggregate and resolves its
// exclusive prefix by decoupled look-back over the preceding tiles' status words; survivors are written in entry order.
// status word: bits 31..30 = 0 invalid / 1 aggregate / 2 inclusive prefix, bits 29..0 = count (entry tables are < 2^24).
constexpr int kCompactThreads = 256;
constexpr int kCompactItems = 32; // per thread: two 16-byte loads, one 32-bit survivor mask
constexpr int kCompactTile = kCompactThreads * kCompactItems;

struct CompactDev {
	const uint8_t *vis;
	uint32_t n;
	uint32_t bit;             // which view bit
	uint32_t *out;            // surviving entry indices, in order
	uint32_t *out2;           // nullable: second copy, each index + add2 — this rank's region of the gather destination's inbox
	uint32_t add2;            //   (peer memory over NVLink, hgb_kernels_peer.cuh): the result leaves the GPU from the kernel that produces it
	uint32_t out2_half;       //   words between the two inbox halves; the half is the parity of *out2_step (the gather's step counter)
	const uint32_t *out2_step;
	uint32_t *count;          // total survivors
	uint32_t *ticket;         // zeroed before launch
	volatile uint32_t *status; // n_tiles words, zeroed before launch
};

__global__ void __launch_bounds__(kCompactThreads) k_compact(CompactDev c) {
	__shared__ uint32_t s_tile, s_warp_sum[kCompactThreads / 32], s_prefix;
	__shared__ uint32_t s_list[kCompactTile];
	// chained to the last level kernel (or the previous compaction) by a programmatic dependency: resident early, held here until the
	// visibility bytes are complete (both are no-ops in a plain launch)
	asm volatile("griddepcontrol.launch_dependents;");
	asm volatile("griddepcontrol.wait;" ::: "memory");
	// A CTA takes tiles by ticket until none is left. One tile per CTA (grid = tiles) is the fastest form when the compaction is on the
	// frame's critical path; as the asynchronous tail of the frame it is launched with a FEW persistent CTAs instead (hgb_ctx:
	// compact_ctas), because a full grid of 32 KB CTAs on the high-priority stream takes the shared memory of every SM away from the
	// level kernels of the next frame for as long as its stores to the gather destination are held up by the fabric.
	const uint32_t n_tiles = (c.n + kCompactTile - 1) / kCompactTile;
	for (;;) {
	__syncthreads(); // s_tile / s_list of the previous tile are no longer read
	if (threadIdx.x == 0)
		s_tile = atomicAdd(c.ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_tile;
	if (tile >= n_tiles)
		break;
	const uint32_t base = tile * kCompactTile + threadIdx.x * kCompactItems;

	// 32 flags per thread, two 128-bit loads (vis buffers are padded to a tile multiple and zero-filled); bit `c.bit` of every byte is
	// gathered into one 32-bit mask: (w >> bit) & 0x01010101 leaves one bit per byte, * 0x10204080 moves the four of them to bits 28..31
	const uint4 *src = reinterpret_cast<const uint4 *>(c.vis + base);
	const uint4 ra = src[0], rb = src[1];
	const uint32_t w[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
	uint32_t mask = 0;
#pragma unroll
	for (int i = 0; i < 8; ++i)
		mask |= ((((w[i] >> c.bit) & 0x01010101u) * 0x10204080u) >> 28) << (4 * i);
	if (base + kCompactItems > c.n) { // ragged tail: drop bits past n
		const uint32_t valid = base < c.n ? c.n - base : 0;
		mask &= valid >= 32 ? 0xffffffffu : (1u << valid) - 1u;
	}
	const uint32_t cnt = __popc(mask);

	// block-wide exclusive scan of cnt
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t incl = cnt;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
		if (lane >= d)
			incl += v;
	}
	if (lane == 31)
		s_warp_sum[warp] = incl;
	__syncthreads();
	uint32_t warp_off = 0, tile_total = 0;
#pragma unroll
	for (int w = 0; w < kCompactThreads / 32; ++w) {
		const uint32_t v = s_warp_sum[w];
		if (w < warp)
			warp_off += v;
		tile_total += v;
	}

	// decoupled look-back by warp 0
	if (warp == 0) {
		if (lane == 0) {
			c.status[tile] = (tile == 0 ? 2u << 30 : 1u << 30) | tile_total;
			__threadfence();
		}
		uint32_t prefix = 0;
		if (tile > 0) {
			int look = int(tile) - 1 - int(lane);
			while (true) {
				uint32_t st = look >= 0 ? c.status[look] : (2u << 30); // before tile 0: inclusive prefix 0
				while (__any_sync(0xffffffffu, (st >> 30) == 0)) {      // wait until the whole window is published
					if ((st >> 30) == 0)
						st = c.status[look];
				}
				const uint32_t has_incl = __ballot_sync(0xffffffffu, (st >> 30) == 2);
				// sum aggregates of the lanes closer than the first inclusive prefix (lane order = distance)
				const int first = has_incl ? __ffs(has_incl) - 1 : 32;
				uint32_t contrib = (int(lane) <= first) ? (st & 0x3fffffffu) : 0u;
#pragma unroll
				for (int d = 16; d > 0; d >>= 1)
					contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
				prefix += contrib;
				if (has_incl)
					break;
				look -= 32;
			}
			if (lane == 0) {
				c.status[tile] = (2u << 30) | (prefix + tile_total);
				__threadfence();
			}
		}
		if (lane == 0) {
			s_prefix = prefix;
			if (uint64_t(tile + 1) * kCompactTile >= c.n)
				*c.count = prefix + tile_total; // last tile owns the grand total
		}
	}
	__syncthreads();

	// survivors of this tile, ascending, staged in shared memory and written out as one contiguous run: full-line coalesced stores
	// whatever the visible fraction — and the only form in which the copy for the gather destination (out2, peer memory over NVLink)
	// travels well: per-thread 4-byte stores cost one fabric transaction each
	uint32_t r = warp_off + (incl - cnt);
	while (mask) {
		s_list[r++] = base + (__ffs(mask) - 1);
		mask &= mask - 1;
	}
	__syncthreads();
	const uint32_t prefix = s_prefix;
	if (c.out2) {
		uint32_t *out2 = c.out2 + size_t(*c.out2_step & 1u) * c.out2_half;
		for (uint32_t i = threadIdx.x; i < tile_total; i += kCompactThreads) {
			const uint32_t e = s_list[i];
			c.out[prefix + i] = e;
			out2[prefix + i] = e + c.add2;
		}
	} else {
		for (uint32_t i = threadIdx.x; i < tile_total; i += kCompactThreads)
			c.out[prefix + i] = s_list[i];
	}
	} // next tile
}

// ---- all views of an entry table in ONE pass -----------------------------------------------------------------------------------------
// The forward pipeline culls the same opaque list against up to 6 frusta per frame (view + 4 PSSM splits + spot,
// engine/forward_pipeline.cpp:331-332,380-381). One launch per view re-reads the visibility bytes and walks its own look-back chain, one
// after the other; here a tile's bytes are read ONCE, every view's bit is peeled off the same registers, and the NV look-back chains
// advance side by side — warp v resolves view v (8 warps, up to 8 views). Output lists, counts and status words are per view, laid out
// like NV independent k_compact launches would leave them (k_segment_ranks and hgb_get_visible read them unchanged).
struct CompactMultiDev {
	const uint8_t *vis;
	uint32_t n;
	uint32_t *out[kMaxViews];  // surviving entry indices per view, in order
	uint32_t *count;           // kMaxViews totals
	uint32_t *ticket;          // one word, zeroed before launch
	volatile uint32_t *status; // [view][n_tiles] words, zeroed before launch
	uint32_t n_tiles;
};

template <int NV> __global__ void __launch_bounds__(kCompactThreads) k_compact_multi(CompactMultiDev c) {
	__shared__ uint32_t s_tile, s_warp_sum[NV][kCompactThreads / 32], s_prefix[NV], s_total[NV];
	__shared__ uint32_t s_list[kCompactTile];
	asm volatile("griddepcontrol.launch_dependents;");
	asm volatile("griddepcontrol.wait;" ::: "memory");
	if (threadIdx.x == 0)
		s_tile = atomicAdd(c.ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_tile;
	const uint32_t base = tile * kCompactTile + threadIdx.x * kCompactItems;
	const uint4 *src = reinterpret_cast<const uint4 *>(c.vis + base);
	const uint4 ra = src[0], rb = src[1];
	const uint32_t w[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
	uint32_t valid_mask = 0xffffffffu;
	if (base + kCompactItems > c.n) { // ragged tail: drop bits past n
		const uint32_t valid = base < c.n ? c.n - base : 0;
		valid_mask = valid >= 32 ? 0xffffffffu : (1u << valid) - 1u;
	}
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t mask[NV], excl[NV];
#pragma unroll
	for (int v = 0; v < NV; ++v) {
		uint32_t m = 0;
#pragma unroll
		for (int i = 0; i < 8; ++i)
			m |= ((((w[i] >> v) & 0x01010101u) * 0x10204080u) >> 28) << (4 * i);
		m &= valid_mask;
		mask[v] = m;
		const uint32_t cnt = __popc(m);
		uint32_t incl = cnt;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
			if (lane >= d)
				incl += t;
		}
		if (lane == 31)
			s_warp_sum[v][warp] = incl;
		excl[v] = incl - cnt;
	}
	__syncthreads();
	// warp v: decoupled look-back of view v
	if (warp < NV) {
		const int v = int(warp);
		volatile uint32_t *status = c.status + size_t(v) * c.n_tiles;
		uint32_t tile_total = 0;
#pragma unroll
		for (int k = 0; k < kCompactThreads / 32; ++k)
			tile_total += s_warp_sum[v][k];
		if (lane == 0) {
			status[tile] = (tile == 0 ? 2u << 30 : 1u << 30) | tile_total;
			__threadfence();
		}
		uint32_t prefix = 0;
		if (tile > 0) {
			int look = int(tile) - 1 - int(lane);
			while (true) {
				uint32_t st = look >= 0 ? status[look] : (2u << 30);
				while (__any_sync(0xffffffffu, (st >> 30) == 0)) {
					if ((st >> 30) == 0)
						st = status[look];
				}
				const uint32_t has_incl = __ballot_sync(0xffffffffu, (st >> 30) == 2);
				const int first = has_incl ? __ffs(has_incl) - 1 : 32;
				uint32_t contrib = (int(lane) <= first) ? (st & 0x3fffffffu) : 0u;
#pragma unroll
				for (int d = 16; d > 0; d >>= 1)
					contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
				prefix += contrib;
				if (has_incl)
					break;
				look -= 32;
			}
			if (lane == 0) {
				status[tile] = (2u << 30) | (prefix + tile_total);
				__threadfence();
			}
		}
		if (lane == 0) {
			s_prefix[v] = prefix;
			s_total[v] = tile_total;
			if (uint64_t(tile + 1) * kCompactTile >= c.n)
				c.count[v] = prefix + tile_total;
		}
	}
	__syncthreads();
	// survivors, one view after the other through the same staging buffer
#pragma unroll
	for (int v = 0; v < NV; ++v) {
		uint32_t warp_off = 0;
#pragma unroll
		for (int k = 0; k < kCompactThreads / 32; ++k)
			warp_off += k < int(warp) ? s_warp_sum[v][k] : 0u;
		uint32_t r = warp_off + excl[v], m = mask[v];
		while (m) {
			s_list[r++] = base + (__ffs(m) - 1);
			m &= m - 1;
		}
		__syncthreads();
		const uint32_t total = s_total[v], prefix = s_prefix[v];
		uint32_t *out = c.out[v];
		for (uint32_t i = threadIdx.x; i < total; i += kCompactThreads)
			out[prefix + i] = s_list[i];
		__syncthreads();
	}
}

// Batched scene instances (one camera per instance): the global stable compaction already keeps each instance's survivors
// contiguous; the exclusive rank of every instance's first entry gives the per-instance offsets/counts. One warp per
// instance: inclusive prefix of the preceding tile (left in status[] by k_compact) + popcount of the flags between the tile
// start and the instance start.
__global__ void __launch_bounds__(256) k_segment_ranks(const uint8_t *__restrict__ vis, uint32_t bit, const uint32_t *__restrict__ seg_start, uint32_t n_seg,
	const volatile uint32_t *status, uint32_t *__restrict__ seg_rank) {
	const uint32_t seg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
	asm volatile("griddepcontrol.wait;" ::: "memory"); // see k_compact
	if (seg >= n_seg)
		return;
	const uint32_t e = seg_start[seg];
	const uint32_t tile = e / kCompactTile;
	uint32_t cnt = 0;
	for (uint32_t i = tile * kCompactTile + lane; i < e; i += 32)
		cnt += (vis[i] >> bit) & 1u;
#pragma unroll
	for (int d = 16; d > 0; d >>= 1)
		cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
	if (lane == 0)
		seg_rank[seg] = cnt + (tile > 0 ? (status[tile - 1] & 0x3fffffffu) : 0u);
}

} // namespace hgb
