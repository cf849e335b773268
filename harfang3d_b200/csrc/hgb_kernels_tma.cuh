// Persistent, TMA-staged form of the level kernel (OPT-IN A/B variant, hgb_options.tma_level: measured 4 % slower than the default
// one-thread-per-node kernel on the 16M-node batch, profiles/README.md r01c). Same arithmetic and outputs as k_update_level (hgb_kernels.cuh); what
// changes is how bytes move:
//   * a node tile is 1536 contiguous bytes (12 rows x 32 positions), so one lane fetches it with ONE bulk copy
//     (cp.async.bulk global -> shared, completion on an mbarrier) instead of twelve 128-byte LDGs per warp plus their address math;
//   * warps are persistent (grid = resident CTAs of the device) and walk the level's tiles with a two-stage ring per warp: the
//     copy of the warp's NEXT tile is in flight while the current one is computed, so the DRAM latency of the tile stream is
//     never exposed and no L2 look-ahead prefetch instructions are needed;
//   * the 32 world matrices of a tile (again 1536 contiguous bytes) are written back into the stage buffer and leave the SM as one
//     bulk store (shared -> global); only tiles that straddle a level boundary use masked stores.
#pragma once
#include "hgb_kernels.cuh"

namespace hgb {

constexpr uint32_t kTileBytes = kTileWords * 4; // 1536

__device__ __forceinline__ void bulk_store(void *dst, uint32_t src_smem, uint32_t bytes) {
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
	asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int kTmaWarps = 8;

template <bool FUSED, bool PER_INSTANCE>
__global__ void __launch_bounds__(kTmaWarps * 32, 6) k_update_level_tma(SceneDev s, Views views, uint32_t begin, uint32_t end) {
	__shared__ __align__(128) uint32_t s_tile[kTmaWarps][2][kTileWords]; // per warp: two stages; a stage is reused as the world-row staging
	__shared__ __align__(8) uint64_t s_bar[kTmaWarps][2];
	__shared__ float4 s_planes[kMaxViews * 12];

	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	if (threadIdx.x == 0) {
#pragma unroll
		for (int w = 0; w < kTmaWarps; ++w)
			mbar_init(smem_u32(&s_bar[w][0]), 1), mbar_init(smem_u32(&s_bar[w][1]), 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	if (FUSED && !PER_INSTANCE)
		stage_planes(s_planes, views);
	__syncthreads();

	const uint32_t tile_last = (end - 1u) >> 5;
	const uint32_t stride = gridDim.x * kTmaWarps;
	uint32_t tile = (begin >> 5) + blockIdx.x * kTmaWarps + warp;
	const uint32_t bar0 = smem_u32(&s_bar[warp][0]);
	const uint32_t buf0 = smem_u32(&s_tile[warp][0][0]);
	const uint8_t *tiles_b = reinterpret_cast<const uint8_t *>(s.tiles);

	if (tile <= tile_last && lane == 0) {
		mbar_arrive_expect_tx(bar0, kTileBytes);
		bulk_load(buf0, tiles_b + size_t(tile) * kTileBytes, kTileBytes, bar0);
	}
	for (uint32_t it = 0; tile <= tile_last; tile += stride, ++it) {
		const uint32_t st = it & 1u;
		const uint32_t next = tile + stride;
		if (next <= tile_last && lane == 0) { // the other stage was released at the end of the previous iteration
			const uint32_t bar_n = bar0 + (st ^ 1u) * 8u;
			mbar_arrive_expect_tx(bar_n, kTileBytes);
			bulk_load(buf0 + (st ^ 1u) * kTileBytes, tiles_b + size_t(next) * kTileBytes, kTileBytes, bar_n);
		}
		const uint32_t *t = &s_tile[warp][st][lane];
		const uint32_t p = (tile << 5) + lane;
		const bool live = p >= begin && p < end;
		while (!mbar_try_wait(bar0 + st * 8u, (it >> 1) & 1u)) {
		}

		Mat4 world;
		uint32_t meta = t[kRowEntMeta * 32], eidx = t[kRowEntIdx * 32];
		{
			float px = __uint_as_float(t[0]), py = __uint_as_float(t[32]), pz = __uint_as_float(t[64]);
			float ex = __uint_as_float(t[96]), ey = __uint_as_float(t[128]), ez = __uint_as_float(t[160]);
			float sx = __uint_as_float(t[192]), sy = __uint_as_float(t[224]), sz = __uint_as_float(t[256]);
			const uint32_t par = live ? t[kRowParent * 32] : kRoot;
			if (!live)
				meta = 0;
			const bool preset = live && s.preset && s.preset[p];
			if (par != kRoot && !preset) { // start the parent row on its way while the trigonometry runs (no registers held)
				prefetch_l1(s.world + size_t(par) * 12);
				prefetch_l1(s.world + size_t(par) * 12 + 11);
			}
			if (meta & kLaneAnimatedMask) { // track-driven triples: fetch the value sampled this frame (or the last one sampled)
				if (meta & kLaneAnimated) {
					const float4 v = s.anim_val[__float_as_uint(px)];
					px = v.x, py = v.y, pz = v.z;
				}
				if (meta & (kLaneAnimated << 1)) {
					const float4 v = s.anim_val[__float_as_uint(ex)];
					ex = v.x, ey = v.y, ez = v.z;
				}
				if (meta & (kLaneAnimated << 2)) {
					const float4 v = s.anim_val[__float_as_uint(sx)];
					sx = v.x, sy = v.y, sz = v.z;
				}
			}
			if (preset) {
				world = load_mat4(s.world + size_t(p) * 12);
			} else {
				world = transformation_mat4(px, py, pz, ex, ey, ez, sx, sy, sz);
				if (par != kRoot)
					world = mat4_mul(load_mat4(s.world + size_t(par) * 12), world);
			}
		}

		// world rows: AoS into the stage buffer (every lane has read its inputs), then one bulk store for a full tile
		__syncwarp();
		float4 *so = reinterpret_cast<float4 *>(&s_tile[warp][st][0]);
		so[lane * 3 + 0] = make_float4(world.m[0][0], world.m[0][1], world.m[0][2], world.m[0][3]);
		so[lane * 3 + 1] = make_float4(world.m[1][0], world.m[1][1], world.m[1][2], world.m[1][3]);
		so[lane * 3 + 2] = make_float4(world.m[2][0], world.m[2][1], world.m[2][2], world.m[2][3]);
		const bool full = (tile << 5) >= begin && (tile << 5) + 32u <= end; // warp-uniform
		float4 *dst = reinterpret_cast<float4 *>(s.world) + size_t(tile) * 96;
		if (full) {
			fence_proxy_async();
			__syncwarp();
			if (lane == 0)
				bulk_store(dst, buf0 + st * kTileBytes, kTileBytes);
		} else { // tile straddles a level boundary: only this level's rows may be written
			__syncwarp();
			const uint32_t lo = max(begin, tile << 5) - (tile << 5), hi = min(end, (tile << 5) + 32u) - (tile << 5);
#pragma unroll
			for (int j = 0; j < 3; ++j) {
				const uint32_t k = lane + 32u * j;
				if (k >= lo * 3u && k < hi * 3u)
					dst[k] = so[k];
			}
		}

		if (FUSED && (meta & kEntHas)) {
			const uint32_t bits = entry_visibility<PER_INSTANCE>(s, world, meta, s_planes, views.n_views, p);
			((meta & kEntKindTransparent) ? s.vis1 : s.vis0)[eidx] = uint8_t(bits);
			if (meta & kEntMore) {
				const uint32_t e0 = s.ent_off[p], e1 = s.ent_off[p + 1];
				for (uint32_t k = e0; k < e1; ++k) {
					const uint2 rec = s.pent[k];
					const uint32_t b = entry_visibility<PER_INSTANCE>(s, world, rec.y, s_planes, views.n_views, p);
					((rec.x >> 31) ? s.vis1 : s.vis0)[rec.x & 0x7fffffffu] = uint8_t(b);
				}
			}
		}

		// release the stage: the bulk store must have read it, and every lane's generic accesses must be ordered before the next
		// bulk load (async proxy) that lane 0 issues into it at the top of the next iteration
		if (lane == 0) {
			if (full)
				bulk_store_wait_read();
			else
				fence_proxy_async();
		}
		__syncwarp();
	}
}

} // namespace hgb
