// hgb_compact_visible_noalloc: the order-preserving compaction of a visibility-byte array (k_compact, hgb_kernels.cuh) for callers that
// cannot allocate — code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc. No counterpart in the
// reference. Nothing in this file calls cudaMalloc, cudaFree or an allocating library
// routine: inputs and outputs are the caller's device buffers and the kernel's scratch — the tile ticket and the look-back status
// words that the context path keeps in per-table device buffers — lives in ONE file-scope __device__ array of a fixed maximum size.
//
// That array is held by every GPU that loads the library until the process exits (8.25 KB). Every call on a device shares it and it
// keeps the last call's contents: each call clears the words it is about to use on the caller's stream, ahead of its kernel. Calls
// whose GPU work overlaps (two streams, or two threads on one device) RACE on it — serialise them.
//
// In host code the array's name is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, taken here once per device with that device current.
#include "../../include/hgb200.h"

#include <cuda_runtime.h>
#include <mutex>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

extern "C" __attribute__((visibility("hidden"))) void hgbi_set_error(const char *msg); // hgb_ctx.cu: what hgb_last_error() returns

namespace {

constexpr int kThreads = 256, kItems = 32, kTile = kThreads * kItems; // the tiling of k_compact: 8192 entries per CTA
constexpr uint32_t kMaxEntries = HGB_NOALLOC_MAX_ENTRIES;
constexpr uint32_t kMaxTiles = (kMaxEntries + kTile - 1) / kTile; // 2048
constexpr uint32_t kStatusOff = 64;                               // words: [0] tile ticket, [64 ...] one status word per tile
constexpr int kMaxDevices = 64;

static __device__ __align__(256) uint32_t g_compact_scratch[kStatusOff + kMaxTiles];

int fail(const char *fmt, ...) {
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	hgbi_set_error(buf);
	return 1;
}

// Same algorithm as k_compact (single pass, tiles taken in ticket order, decoupled look-back, survivors staged in shared memory and
// written as one contiguous run per tile). Differences: the scratch is a parameter (the static array), the visibility bytes need no
// padding to a tile multiple (the ragged tail is read byte by byte), and writes stop at the caller's capacity.
__global__ void __launch_bounds__(kThreads) k_compact_static(const uint8_t *__restrict__ vis, uint32_t n, uint32_t bit, uint32_t *__restrict__ out, uint32_t capacity,
	uint32_t *__restrict__ count, uint32_t *scratch) {
	__shared__ uint32_t s_tile, s_warp_sum[kThreads / 32], s_prefix;
	__shared__ uint32_t s_list[kTile];
	uint32_t *ticket = scratch;
	volatile uint32_t *status = scratch + kStatusOff;
	if (threadIdx.x == 0)
		s_tile = atomicAdd(ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_tile;
	const uint32_t base = tile * kTile + threadIdx.x * kItems;

	uint32_t mask = 0;
	if (base + kItems <= n) { // 32 flags: two 128-bit loads; bit `bit` of every byte gathered into one 32-bit mask
		const uint4 *src = reinterpret_cast<const uint4 *>(vis + base);
		const uint4 ra = src[0], rb = src[1];
		const uint32_t w[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
#pragma unroll
		for (int i = 0; i < 8; ++i)
			mask |= ((((w[i] >> bit) & 0x01010101u) * 0x10204080u) >> 28) << (4 * i);
	} else {
		for (uint32_t i = 0; i < uint32_t(kItems) && base + i < n; ++i)
			mask |= uint32_t((vis[base + i] >> bit) & 1u) << i;
	}
	const uint32_t cnt = __popc(mask);

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
	for (int w = 0; w < kThreads / 32; ++w) {
		const uint32_t v = s_warp_sum[w];
		if (w < int(warp))
			warp_off += v;
		tile_total += v;
	}

	if (warp == 0) { // decoupled look-back: status = flag (1 aggregate, 2 inclusive prefix) << 30 | value
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
			s_prefix = prefix;
			if (uint64_t(tile + 1) * kTile >= n)
				*count = prefix + tile_total; // the last tile owns the grand total (may exceed the capacity: nothing past it is written)
		}
	}
	__syncthreads();

	uint32_t r = warp_off + (incl - cnt);
	while (mask) {
		s_list[r++] = base + (__ffs(mask) - 1);
		mask &= mask - 1;
	}
	__syncthreads();
	const uint32_t prefix = s_prefix;
	for (uint32_t i = threadIdx.x; i < tile_total; i += kThreads)
		if (prefix + i < capacity)
			out[prefix + i] = s_list[i];
}

std::mutex g_setup_mutex;
uint32_t *g_scratch_of_device[kMaxDevices] = {}; // device address of g_compact_scratch on each device that has asked for it

} // namespace

extern "C" int hgb_compact_visible_noalloc(int device, const uint8_t *d_vis, uint32_t n, uint32_t view_bit, uint32_t *d_out, uint32_t out_capacity, uint32_t *d_count,
	void *cuda_stream) {
	if (n > kMaxEntries)
		return fail("hgb_compact_visible_noalloc: %u entries exceed the maximum of %u this entry point's static scratch covers (HGB_NOALLOC_MAX_ENTRIES); "
					"use a context (hgb_update / hgb_cull + hgb_get_visible) for larger tables",
			n, kMaxEntries);
	if (view_bit > 7)
		return fail("hgb_compact_visible_noalloc: view bit %u (a visibility byte holds views 0..7)", view_bit);
	if (!d_count || (n && (!d_vis || (!d_out && out_capacity))))
		return fail("hgb_compact_visible_noalloc: null buffer");
	if (reinterpret_cast<uintptr_t>(d_vis) & 15u)
		return fail("hgb_compact_visible_noalloc: the visibility bytes must be 16-byte aligned");
	if (device < 0 || device >= kMaxDevices)
		return fail("hgb_compact_visible_noalloc: device %d out of range", device);
	cudaError_t e = cudaSetDevice(device);
	if (e != cudaSuccess)
		return fail("hgb_compact_visible_noalloc: cudaSetDevice(%d): %s (there is no CPU path)", device, cudaGetErrorString(e));
	uint32_t *scratch = nullptr;
	{ // set-up, once per device, with that device current: the array's name is only a stand-in on the host
		std::lock_guard<std::mutex> lock(g_setup_mutex);
		if (!g_scratch_of_device[device]) {
			void *p = nullptr;
			e = cudaGetSymbolAddress(&p, g_compact_scratch);
			if (e != cudaSuccess || !p)
				return fail("hgb_compact_visible_noalloc: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
			g_scratch_of_device[device] = static_cast<uint32_t *>(p);
		}
		scratch = g_scratch_of_device[device];
	}
	const cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	if (!n) {
		e = cudaMemsetAsync(d_count, 0, 4, stream);
		return e == cudaSuccess ? 0 : fail("hgb_compact_visible_noalloc: %s", cudaGetErrorString(e));
	}
	const uint32_t n_tiles = (n + kTile - 1) / kTile;
	// the array keeps the last call's ticket and status words: what this call uses starts at zero
	e = cudaMemsetAsync(scratch, 0, size_t(kStatusOff + n_tiles) * 4, stream);
	if (e != cudaSuccess)
		return fail("hgb_compact_visible_noalloc: %s", cudaGetErrorString(e));
	k_compact_static<<<n_tiles, kThreads, 0, stream>>>(d_vis, n, view_bit, d_out, out_capacity, d_count, scratch);
	e = cudaGetLastError();
	if (e != cudaSuccess)
		return fail("hgb_compact_visible_noalloc: launch: %s", cudaGetErrorString(e));
	return 0;
}
