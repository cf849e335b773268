// sm_100a kernels of the scene-update path. All of them are HBM-bound streaming / gather / compaction kernels (no tensor
// cores: ~1.3 flop/B). Data layout: everything on the device is stored in DEPTH-SORTED POSITION order
// ("pos"): position p of a transform slot is its rank in a level-by-level (BFS) ordering with siblings adjacent, so a level
// is a contiguous range, every parent has a smaller position than its children, and sibling threads hit the same parent lines.
#pragma once
#include "hgb_math.cuh"

namespace hgb {

constexpr uint32_t kRoot = 0xffffffffu;
constexpr int kMaxViews = 8;

// entry meta word: bounds index in the low 24 bits, flags in the high 8
constexpr uint32_t kEntDisabled = 1u << 24;
constexpr uint32_t kEntKindTransparent = 1u << 31; // in PosEntry::entry (bit 31): which of the two entry tables

// View frusta live in device memory (6 float4 per view: Top,Bottom,Left,Right,Near,Far, foundation/frustum.h:14; w holds -2*w) so that a camera
// move is a 96-byte copy and never invalidates the captured frame graph.
struct Views {
	const float4 *planes;
	int n_views;
};

struct SceneDev {
	// per position (sorted order)
	const float *trs;           // SoA: 9 arrays of stride trs_stride: px py pz rx ry rz sx sy sz
	size_t trs_stride;
	const uint32_t *parent_pos; // kRoot for roots
	float *world;               // AoS 12 floats per position, 16-B aligned (== hg::Mat4 bit layout)
	const uint8_t *preset;      // nullable: 1 -> world[p] was pre-set this frame (Scene::SetNodeWorldMatrix), keep it
	// fused cull: display entries attached to each position (CSR)
	const uint32_t *ent_off;    // n_pos + 1
	const uint2 *pent;          // {entry index | kind bit, meta}
	const float *bounds6;       // MinMax table, 6 floats per (model, list)
	uint8_t *vis0, *vis1;       // per entry table (opaque, transparent): bit v set -> survives view v
	const float4 *inst_planes;  // per-instance frusta (6 float4 each) or null
	const uint16_t *pent_view;  // per pent record: which instance frustum (only with inst_planes)
};

// ---- world-matrix propagation, one level per launch ---------------------------------------------------------------------
// Replaces Scene::ComputeTransformWorldMatrix (engine/scene.cpp:263-277) for the positions [begin, end) of one hierarchy
// level: local = TransformationMat4(pos, rot, scl); world = parent ? world[parent] * local : local. Parents were written
// by the previous launch. With FUSED the freshly computed world matrix is tested, still in registers, against the
// view frusta for every display entry attached to the position (GetModelDisplayLists flag logic + CullModelDisplayLists,
// engine/scene.cpp:330-364, engine/render_pipeline.cpp:2098-2104) and only a 1-byte visibility mask per entry is written.
template <bool FUSED, bool PER_INSTANCE>
__global__ void __launch_bounds__(256) k_update_level(SceneDev s, Views views, uint32_t begin, uint32_t end) {
	__shared__ float4 s_planes[kMaxViews * 6];
	if (FUSED && !PER_INSTANCE) { // the shared view frusta: one broadcast LDS.128 per plane instead of a global load
		if (threadIdx.x < uint32_t(views.n_views) * 6u)
			s_planes[threadIdx.x] = views.planes[threadIdx.x];
		__syncthreads();
	}
	const uint32_t p = begin + blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= end)
		return;

	Mat4 world;
	if (s.preset && s.preset[p]) {
		world = load_mat4(s.world + size_t(p) * 12);
	} else {
		const float *t = s.trs + p;
		const size_t st = s.trs_stride;
		const float px = t[0], py = t[st], pz = t[2 * st], ex = t[3 * st], ey = t[4 * st], ez = t[5 * st], sx = t[6 * st], sy = t[7 * st], sz = t[8 * st];
		const uint32_t par = s.parent_pos[p];
		world = transformation_mat4(px, py, pz, ex, ey, ez, sx, sy, sz);
		if (par != kRoot)
			world = mat4_mul(load_mat4(s.world + size_t(par) * 12), world);
		store_mat4(s.world + size_t(p) * 12, world);
	}

	if (FUSED) {
		const uint32_t e0 = s.ent_off[p], e1 = s.ent_off[p + 1];
		for (uint32_t k = e0; k < e1; ++k) {
			const uint2 rec = s.pent[k];
			uint32_t bits = 0;
			if (!(rec.y & kEntDisabled)) {
				const Aabb bb = minmax_transform(world, s.bounds6 + size_t(rec.y & 0xffffffu) * 6);
				if (PER_INSTANCE) {
					bits = survives_frustum(s.inst_planes + size_t(s.pent_view[k]) * 6, bb) ? 1u : 0u;
				} else {
					for (int v = 0; v < views.n_views; ++v)
						bits |= survives_frustum(s_planes + v * 6, bb) ? (1u << v) : 0u;
				}
			}
			((rec.x >> 31) ? s.vis1 : s.vis0)[rec.x & 0x7fffffffu] = uint8_t(bits);
		}
	}
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
// tiles are started in order), ballots/popcounts its 2048 visibility bits, publishes its aggregate and resolves its
// exclusive prefix by decoupled look-back over the preceding tiles' status words; survivors are written in entry order.
// status word: bits 31..30 = 0 invalid / 1 aggregate / 2 inclusive prefix, bits 29..0 = count (entry tables are < 2^24).
constexpr int kCompactThreads = 256;
constexpr int kCompactItems = 8;
constexpr int kCompactTile = kCompactThreads * kCompactItems;

struct CompactDev {
	const uint8_t *vis;
	uint32_t n;
	uint32_t bit;             // which view bit
	uint32_t *out;            // surviving entry indices, in order
	uint32_t *count;          // total survivors
	uint32_t *ticket;         // zeroed before launch
	volatile uint32_t *status; // n_tiles words, zeroed before launch
};

__global__ void __launch_bounds__(kCompactThreads) k_compact(CompactDev c) {
	__shared__ uint32_t s_tile, s_warp_sum[kCompactThreads / 32], s_prefix;
	if (threadIdx.x == 0)
		s_tile = atomicAdd(c.ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_tile;
	const uint32_t base = tile * kCompactTile + threadIdx.x * kCompactItems;

	// 8 flags per thread, one 64-bit load (vis buffers are padded to a tile multiple and zero-filled)
	const uint2 raw = *reinterpret_cast<const uint2 *>(c.vis + base);
	uint32_t mask = 0;
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		mask |= ((raw.x >> (8 * i + c.bit)) & 1u) << i;
		mask |= ((raw.y >> (8 * i + c.bit)) & 1u) << (4 + i);
	}
	if (base + kCompactItems > c.n) { // ragged tail: drop bits past n
		const uint32_t valid = base < c.n ? c.n - base : 0;
		mask &= (1u << valid) - 1u;
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

	uint32_t rank = s_prefix + warp_off + (incl - cnt);
#pragma unroll
	for (int i = 0; i < kCompactItems; ++i)
		if (mask & (1u << i))
			c.out[rank++] = base + i;
}

// Batched scene instances (one camera per instance): the global stable compaction already keeps each instance's survivors
// contiguous; the exclusive rank of every instance's first entry gives the per-instance offsets/counts. One warp per
// instance: inclusive prefix of the preceding tile (left in status[] by k_compact) + popcount of the flags between the tile
// start and the instance start.
__global__ void __launch_bounds__(256) k_segment_ranks(const uint8_t *__restrict__ vis, uint32_t bit, const uint32_t *__restrict__ seg_start, uint32_t n_seg,
	const volatile uint32_t *status, uint32_t *__restrict__ seg_rank) {
	const uint32_t seg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
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
