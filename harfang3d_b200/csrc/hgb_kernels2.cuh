// Animation key sampling, skin bone matrices and the layout shuffles around host uploads/downloads.
#pragma once
#include "hgb_kernels.cuh"
#include "hgb_atan2_exact.h"

namespace hgb {

// ---- animation key sampling -----------------------------------------------------------------------------------------
// One thread per bound track. Replaces Scene::EvaluateBoundAnim(BoundToNodeAnim) (engine/scene.cpp:2154-2181) for the
// transform targets: Evaluate(track, t, value) -> TRS lane of the bound node's transform.
//   GetIntervalKeys (engine/animation.h:109-128) is a linear scan for the first key with key.t > t; the binary search
//   below returns the same index (upper_bound) because keys are sorted by time.
struct alignas(16) Track { // 16 B, read as one uint4
	uint32_t out_slot;   // anim_val[] slot of the (node, pos|rot|scale) triple this track drives
	uint32_t key_off;    // first interval record of the track, in 64-byte units (k_build_key_records lays the records out)
	uint32_t key_count;
	uint16_t play;       // index into play_t[]
	uint8_t kind;        // 0 Position, 1 Rotation (Euler vec3), 2 Scale, 3 Rotation (quaternion)
	uint8_t stride;      // 64-byte units between two records of the track: 2 when it shares its 128-byte lines with its neighbour track
};
struct TrackBuild { // bind time: where a track's keys are and where its records go
	uint32_t key_first;  // first key in the compact pool (quaternion keys already offset)
	uint32_t rec_off;    // = Track::key_off
	uint32_t key_count;
	uint8_t kind, stride, pad[2];
};

// One key = one 32-byte record (one DRAM sector), vec3 Hermite and quaternion keys alike:
//   q0 = {t lo, t hi, v.x, v.y}   q1 = {v.z, tension | v.w, bias, -}
// so the four keys of a Hermite evaluation are 128 contiguous bytes and their times come with them (AnimKeyHermiteT<Vec3>,
// AnimKeyT<Quaternion>: engine/animation.h:28-46). Quaternion keys follow the vec3 keys in the same pool.
struct AnimDev {
	const Track *tracks;
	const int64_t *play_t;    // per playing anim: evaluation time (ns), INT64_MIN = not playing
	const uint4 *keys;        // compact pool: 2 x uint4 per key (source of the records)
	const uint4 *rec;         // interval records: 4 x uint4 (64 B, aligned) per key, see k_build_key_records
	uint32_t quat_base;       // first quaternion key in the pool
	uint32_t *cursor;         // per track: lower key of the interval found last frame (search hint only, never changes a result)
	uint4 *tstate;            // per track: {time of key cursor, time of key cursor+1 (INT64_MAX past the end)} — hint, verified against the record
	float4 *anim_val;         // output: one float4 per driven (node, triple); consecutive tracks write consecutive slots
};

__device__ __forceinline__ float time_to_sec_f(int64_t t) { return __double2float_rn(__ddiv_rn(double(t), 1000000000.0)); } // foundation/time.h:13

// ToEuler(Mat3, RO_YXZ) foundation/matrix3.cpp:91-137,141-163 (euler_order_yxz: i=1 j=0 k=2; the odd and the left-handed negations
// cancel). atan2 is the DOUBLE libm function in the reference (float arguments promoted, result narrowed to float). CUDA's double atan2
// (<= 2 ulp) narrows to the same float except when it sits within a few double-ulps of a float rounding boundary; those evaluations
// (~3e-8 of all) are redone in double-double arithmetic and rounded the way a correctly rounded libm rounds (hgb_atan2_exact.h,
// checked on the host against the reference's libm by oracle/exhaustive_atan2.cpp), so the Euler angles are bit-identical.
__device__ __noinline__ float atan2f_ref_slow(float y, float x) {
	const hgb_exact::dd r = hgb_exact::atan2_dd(double(y), double(x));
	return __double2float_rn(r.hi);
}
__device__ __forceinline__ float atan2f_ref(float y, float x) {
	const double lib = atan2(double(y), double(x));
	if (hgb_exact::near_float_midpoint(lib) && x != 0.f && y != 0.f && fabsf(x) <= 3.402823466e+38f && fabsf(y) <= 3.402823466e+38f)
		return atan2f_ref_slow(y, x);
	return __double2float_rn(lib);
}
__device__ __forceinline__ void mat3_to_euler_yxz(float m00, float m02, float m10, float m11, float m12, float m20, float m22, float &ex, float &ey, float &ez) {
	const float s1 = -m12;
	const float c1 = sqrtf(m11 * m11 + m10 * m10);
	float a0, a2;
	const float a1 = atan2f_ref(s1, c1);
	if (c1 > 1.1920929e-07f) {
		a0 = atan2f_ref(m02, m22);
		a2 = atan2f_ref(m10, m11);
	} else {
		a0 = 0.F;
		const float sign_s1 = s1 < 0.F ? -1.F : (s1 > 0.F ? 1.F : 0.F);
		a2 = __double2float_rn(__dmul_rn(double(-sign_s1), double(atan2f_ref(-m20, m00)))); // float * double in the reference: exact, then narrowed
	}
	ex = a1, ey = a0, ez = a2;
}

// ToEuler(Normalize(q)) engine/scene.cpp:2172 : foundation/quaternion.cpp:57-60 (Normalize), :141-166 (ToMatrix3),
// foundation/matrix3.cpp:91-137 (ToEuler, euler_order_yxz: i=1 j=0 k=2), atan2 as above.
__device__ __forceinline__ void quat_to_euler(float x, float y, float z, float w, float &ex, float &ey, float &ez) {
	const float l = sqrtf(x * x + y * y + z * z + w * w);
	if (l != 0.f) {
		x = x / l, y = y / l, z = z / l, w = w / l;
	} else {
		x = y = z = 0.f, w = 1.f;
	}
	const float sqw = w * w, sqx = x * x, sqy = y * y, sqz = z * z;
	const float invs = 1.f / (sqx + sqy + sqz + sqw);
	const float m00 = (sqx - sqy - sqz + sqw) * invs;
	const float m11 = (-sqx + sqy - sqz + sqw) * invs;
	const float m22 = (-sqx - sqy + sqz + sqw) * invs;
	float tmp1 = x * y, tmp2 = z * w;
	const float m10 = 2.f * (tmp1 + tmp2) * invs;
	tmp1 = x * z, tmp2 = y * w;
	const float m20 = 2.f * (tmp1 - tmp2) * invs;
	const float m02 = 2.f * (tmp1 + tmp2) * invs;
	tmp1 = y * z, tmp2 = x * w;
	const float m12 = 2.f * (tmp1 - tmp2) * invs;
	mat3_to_euler_yxz(m00, m02, m10, m11, m12, m20, m22, ex, ey, ez);
}

// Interval records. The sampler's HBM traffic is one gather per track per frame; measured on the B200 (profiles/README.md r01c), a
// window that crosses a 128-byte line costs 70 % more time than an aligned one. So the device keeps, next to the compact pool, one
// 64-BYTE-ALIGNED RECORD PER KEY = everything the evaluation of the interval that starts at key k needs, and nothing else:
//   Hermite vec3 track:  {t(k), t(k+1)} {y1.xyz, y2.x} {y2.yz, m0.xy} {m0.z, m1.xyz}
//   quaternion track:    {t(k), t(k+1)} {q(k)} {q(k+1)} {-}
// y1 / y2 = values of keys k / k+1; m0 / m1 = the two tangents of HermiteInterpolate (foundation/math.h:137-149), which depend on the
// keys k-1 ... k+2 and on tension / bias of key k but NOT on the evaluation time: k_build_key_records computes them once, with the
// very operations (and this TU's -fmad=false) the evaluation would use, so hoisting them changes no bit of any result — and the
// per-frame gather shrinks from four whole keys (128 bytes) to 64. t(k+1) is INT64_MAX in a track's last record.
// DRAM moves 128-byte lines (measured: 64-byte records laid out track after track still cost 128 bytes of DRAM read each, profiles r02),
// so the records of two NEIGHBOUR tracks (consecutive threads of the sampler) are interleaved: record k of the pair's first track is the
// lower half of line k, record k of the second the upper half. Tracks that advance together — the tracks of one anim, instanced anims —
// then share every line they fetch. Every track owns its records (a key range shared by several tracks is laid out once per track).
// Costs 2x the key memory (the 16M-node batch's 64M keys take 4 GB), built on the device at bind time.
constexpr uint32_t kRecU4 = 4; // uint4s per record

// tangents of one component of HermiteInterpolate (foundation/math.h:137-149), in the reference's operation order
__device__ __forceinline__ void hermite_tangents(float y0, float y1, float y2, float y3, float tension, float bias, float &m0, float &m1) {
	m0 = (y1 - y0) * (1.F + bias) * (1.F - tension) / 2.F;
	m0 += (y2 - y1) * (1.F - bias) * (1.F - tension) / 2.F;
	m1 = (y2 - y1) * (1.F + bias) * (1.F - tension) / 2.F;
	m1 += (y3 - y2) * (1.F - bias) * (1.F - tension) / 2.F;
}

__global__ void __launch_bounds__(256) k_build_key_records(const TrackBuild *__restrict__ tracks, uint32_t n_tracks, const uint4 *__restrict__ keys,
	uint4 *__restrict__ rec) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_tracks)
		return;
	const uint32_t n = tracks[i].key_count;
	const bool quat = tracks[i].kind == 3;
	const size_t base = tracks[i].key_first;
	const size_t rec_off = tracks[i].rec_off, stride = tracks[i].stride;
	for (uint32_t k = 0; k < n; ++k) {
		const uint32_t k0 = k > 0 ? k - 1 : 0u, k2 = min(k + 1, n - 1), k3 = min(k + 2, n - 1);
		const uint4 a1 = keys[(base + k) * 2], b1 = keys[(base + k) * 2 + 1];
		const uint4 a2 = keys[(base + k2) * 2], b2 = keys[(base + k2) * 2 + 1];
		uint4 *d = rec + (rec_off + k * stride) * kRecU4;
		d[0] = k + 1 < n ? make_uint4(a1.x, a1.y, a2.x, a2.y) : make_uint4(a1.x, a1.y, 0xffffffffu, 0x7fffffffu);
		if (quat) { // AnimKeyT<Quaternion>: {t, x, y | z, w}
			d[1] = make_uint4(a1.z, a1.w, b1.x, b1.y);
			d[2] = make_uint4(a2.z, a2.w, b2.x, b2.y);
			d[3] = make_uint4(0, 0, 0, 0);
			continue;
		}
		const uint4 a0 = keys[(base + k0) * 2], b0 = keys[(base + k0) * 2 + 1];
		const uint4 a3 = keys[(base + k3) * 2], b3 = keys[(base + k3) * 2 + 1];
		const float tension = __uint_as_float(b1.y), bias = __uint_as_float(b1.z);
		float m0x, m0y, m0z, m1x, m1y, m1z;
		hermite_tangents(__uint_as_float(a0.z), __uint_as_float(a1.z), __uint_as_float(a2.z), __uint_as_float(a3.z), tension, bias, m0x, m1x);
		hermite_tangents(__uint_as_float(a0.w), __uint_as_float(a1.w), __uint_as_float(a2.w), __uint_as_float(a3.w), tension, bias, m0y, m1y);
		hermite_tangents(__uint_as_float(b0.x), __uint_as_float(b1.x), __uint_as_float(b2.x), __uint_as_float(b3.x), tension, bias, m0z, m1z);
		d[1] = make_uint4(a1.z, a1.w, b1.x, a2.z);
		d[2] = make_uint4(a2.w, b2.x, __float_as_uint(m0x), __float_as_uint(m0y));
		d[3] = make_uint4(__float_as_uint(m0z), __float_as_uint(m1x), __float_as_uint(m1y), __float_as_uint(m1z));
	}
}

// time of key k of a track, read from its record
__device__ __forceinline__ int64_t rec_key_time(const uint4 *rec, uint32_t k, uint32_t stride) {
	const uint2 t = *reinterpret_cast<const uint2 *>(rec + size_t(k) * stride * kRecU4);
	return int64_t(uint64_t(t.x) | (uint64_t(t.y) << 32));
}
__device__ __forceinline__ uint32_t upper_bound_rec(const uint4 *rec, uint32_t n, int64_t at, uint32_t stride) {
	uint32_t lo = 0, hi = n;
	while (lo < hi) {
		const uint32_t mid = (lo + hi) >> 1;
		if (rec_key_time(rec, mid, stride) > at)
			hi = mid;
		else
			lo = mid + 1;
	}
	return lo;
}

// One thread per bound track; a warp fetches its 32 records cooperatively — 4 lanes per record, 16 bytes per lane, straight into
// shared memory with cp.async (LDGSTS, L1 bypassed) — so one load instruction covers eight complete records (per-thread 16-byte
// gathers are bound by L1 wavefronts at 89 % of the data-pipe peak, profiles/r01b). Each lane then reads its own record (80-byte
// stride: conflict-free LDS.128). Pass 0 fetches the record the per-track state predicts; when the prediction fails (time jump, loop
// wrap, state never written) pass 1 fetches the right one for the lanes that need it.
constexpr int kSampleWarps = 8;
constexpr uint32_t kWinStride = 80;

__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void *src) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}

__global__ void __launch_bounds__(kSampleWarps * 32, 6) k_sample_tracks(AnimDev a, uint32_t first, uint32_t n_tracks) {
	__shared__ __align__(16) uint8_t s_win[kSampleWarps][32 * kWinStride];
	// the first level kernel may become resident while the last wave of this grid drains (it waits before reading anim_val)
	asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
	bool active = i < first + n_tracks;
	Track trk{};
	int64_t at = INT64_MIN;
	if (active) {
		*reinterpret_cast<uint4 *>(&trk) = *reinterpret_cast<const uint4 *>(a.tracks + i);
		at = a.play_t[trk.play];
		active = at != INT64_MIN && trk.key_count != 0; // not playing / Evaluate() returns false on an empty track (animation.h:156)
	}
	if (!__any_sync(0xffffffffu, active))
		return;
	const uint32_t n = trk.key_count;
	const uint32_t kbase = trk.key_off, kstride = trk.stride;
	const uint4 *rec = a.rec + size_t(kbase) * kRecU4;
	uint8_t *sw = s_win[warp];
	const uint4 *win = reinterpret_cast<const uint4 *>(sw + lane * kWinStride);
	const uint32_t sub = lane & 3u, grp = lane >> 2;
	const uint8_t *pool = reinterpret_cast<const uint8_t *>(a.rec);
	auto t_lo = [](const uint4 &k) { return int64_t(uint64_t(k.x) | (uint64_t(k.y) << 32)); };
	auto t_hi = [](const uint4 &k) { return int64_t(uint64_t(k.z) | (uint64_t(k.w) << 32)); };

	// Which record? The per-track state {time of the cursor's key, time of the next key} (16 bytes, read coalesced) tells BEFORE the
	// fetch whether this frame stays in last frame's interval or enters the next one, so that one aligned record per track is all a
	// frame reads, whether or not the track advances. The state is a hint: the fetched record carries both key times and is verified;
	// anything unexpected (time jump, loop wrap, state never written) falls back to the binary search + refetch.
	uint32_t cur = 0, kf1 = 0, ub = 0;
	int64_t t_cur = 0, t_next = 0;
	bool want = active;
	if (active) {
		cur = min(a.cursor[i], n - 1);
		const uint4 st = a.tstate[i];
		t_cur = t_lo(st), t_next = t_hi(st);
		kf1 = (cur + 1 < n && at >= t_next && t_next > t_cur) ? cur + 1 : cur;
	}
#pragma unroll 1
	for (int pass = 0; pass < 2; ++pass) {
		// cooperative fetch: round r serves tracks 8r..8r+7
		const uint32_t want_rec = want ? kbase + kf1 * kstride : 0xffffffffu;
#pragma unroll
		for (int r = 0; r < 4; ++r) {
			const uint32_t tr = r * 8 + grp;
			const uint32_t k = __shfl_sync(0xffffffffu, want_rec, tr);
			if (k != 0xffffffffu)
				cp_async16(uint32_t(__cvta_generic_to_shared(sw + tr * kWinStride + sub * 16)), pool + size_t(k) * 64 + sub * 16);
		}
		asm volatile("cp.async.wait_all;" ::: "memory");
		__syncwarp();
		if (pass == 1)
			break;
		want = false;
		if (active) {
			// record kf1 covers [key kf1, key kf1+1): the right one when key kf1 <= at < key kf1+1, or at/after the last key (its
			// t(k+1) is INT64_MAX), or before the first one
			const uint4 h = win[0];
			const int64_t tB = t_lo(h), tC = t_hi(h);
			if (tB <= at && at < tC) {
				ub = kf1 + 1;
			} else if (kf1 == 0 && at < tB) {
				ub = 0;
			} else {
				ub = upper_bound_rec(rec, n, at, kstride);
				const uint32_t k = ub > 0 ? ub - 1 : 0; // ub == 0: before the first key -> key 0; ub == n: at/after the last key -> key n-1
				want = k != kf1;
				kf1 = k;
			}
		}
		if (!__any_sync(0xffffffffu, want))
			break;
		__syncwarp(); // every lane has read its pass-0 record before pass 1 overwrites windows
	}
	if (!active)
		return;
	const uint4 h = win[0], r1 = win[1], r2 = win[2];
	{ // the hint of the next frame (written only when it changes: once per interval)
		const int64_t tB = t_lo(h), tC = t_hi(h);
		if (kf1 != cur)
			a.cursor[i] = kf1;
		if (tB != t_cur || tC != t_next)
			a.tstate[i] = h;
	}
	float4 *dst = a.anim_val + trk.out_slot;
	const int64_t t1 = t_lo(h), t2 = t_hi(h);

	if (trk.kind == 3) { // AnimTrackT<Quaternion>: EvaluateLinear animation.h:141-153, then ToEuler(Normalize())
		float4 q = make_float4(__uint_as_float(r1.x), __uint_as_float(r1.y), __uint_as_float(r1.z), __uint_as_float(r1.w));
		if (!(ub == 0 || ub == n)) {
			const float4 q0 = q;
			const float4 q1 = make_float4(__uint_as_float(r2.x), __uint_as_float(r2.y), __uint_as_float(r2.z), __uint_as_float(r2.w));
			const float k = time_to_sec_f(at - t1) / time_to_sec_f(t2 - t1);
			q = make_float4(q0.x + (q1.x - q0.x) * k, q0.y + (q1.y - q0.y) * k, q0.z + (q1.z - q0.z) * k, q0.w + (q1.w - q0.w) * k);
		}
		float ex, ey, ez;
		quat_to_euler(q.x, q.y, q.z, q.w, ex, ey, ez);
		*dst = make_float4(ex, ey, ez, 0.f);
		return;
	}
	// AnimTrackHermiteT<Vec3>: EvaluateHermite animation.h:155-169 with the tangents taken from the record
	float ox = __uint_as_float(r1.x), oy = __uint_as_float(r1.y), oz = __uint_as_float(r1.z); // key kf1 itself: before the first / after the last key
	if (!(ub == 0 || ub == n)) {
		const uint4 r3 = win[3];
		const float u = time_to_sec_f(at - t1) / time_to_sec_f(t2 - t1);
		const float t2_ = u * u, t3 = t2_ * u;
		const float a0 = 2.F * t3 - 3.F * t2_ + 1.F;
		const float a1 = t3 - 2.F * t2_ + u;
		const float a2 = t3 - t2_;
		const float a3 = -2.F * t3 + 3.F * t2_;
		ox = ox * a0 + __uint_as_float(r2.z) * a1 + __uint_as_float(r3.y) * a2 + __uint_as_float(r1.w) * a3;
		oy = oy * a0 + __uint_as_float(r2.w) * a1 + __uint_as_float(r3.z) * a2 + __uint_as_float(r2.x) * a3;
		oz = oz * a0 + __uint_as_float(r3.x) * a1 + __uint_as_float(r3.w) * a2 + __uint_as_float(r2.y) * a3;
	}
	*dst = make_float4(ox, oy, oz, 0.f);
}

// ---- skin bone matrices ----------------------------------------------------------------------------------------------
// _DrawSkinnedModelDisplayLists engine/render_pipeline.cpp:2172-2173 (+ :2177-2181 previous frame): per (skinned list, bone)
// world[mtx_idxs[j]] * bind_pose[bones_idxs[j]] -> to_bgfx (:59-62): 4x4 column-major with last row 0 0 0 1.
struct BoneDev {
	const uint32_t *mtx_pos;  // per bone: position of the bone's transform
	const uint32_t *bind_idx; // per bone: index into bind_pose
	const float *bind_pose;   // 12 floats each
	const float *world;
	const float *world_prev;  // nullable
	float *out;               // 16 floats per bone
	float *out_prev;
	uint32_t n;
};

__device__ __forceinline__ void store_bgfx(float *o, const Mat4 &m) {
	float4 *q = reinterpret_cast<float4 *>(o);
	q[0] = make_float4(m.m[0][0], m.m[1][0], m.m[2][0], 0.f);
	q[1] = make_float4(m.m[0][1], m.m[1][1], m.m[2][1], 0.f);
	q[2] = make_float4(m.m[0][2], m.m[1][2], m.m[2][2], 0.f);
	q[3] = make_float4(m.m[0][3], m.m[1][3], m.m[2][3], 1.f);
}

__global__ void __launch_bounds__(256) k_bone_matrices(BoneDev b) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= b.n)
		return;
	const uint32_t mp = b.mtx_pos[i];
	const Mat4 bind = load_mat4(b.bind_pose + size_t(b.bind_idx[i]) * 12);
	store_bgfx(b.out + size_t(i) * 16, mat4_mul(load_mat4(b.world + size_t(mp) * 12), bind));
	if (b.world_prev)
		store_bgfx(b.out_prev + size_t(i) * 16, mat4_mul(load_mat4(b.world_prev + size_t(mp) * 12), bind));
}

// ---- opt-in culling of skinned display lists ----------------------------------------------------------------------------------------
// The reference never culls skinned lists (engine/scene_forward_pipeline.cpp:560-562, "FIXME cull skinned models !!!"); the default here
// is the same. This is the opt-in: a skinned list survives unless the UNION over its bones of the list's MinMax carried through the
// bone's skinning matrix world[mtx_idxs[j]] * bind_pose[bones_idxs[j]] (render_pipeline.cpp:2172-2173) — a conservative bound of every
// skinned vertex, whatever its weights — is V_Outside (TestVisibility, foundation/frustum.cpp:59-79). One warp per list (<= 32 bones,
// render_pipeline.h:38): lane j builds its bone's box with the path's own operator*(Mat4,Mat4) and operator*(Mat4,MinMax), the boxes
// are joined by shuffles (min / max are exact), lane 0 tests the planes.
struct SkinCullDev {
	const uint32_t *bone_off;  // n_lists + 1: bones of list i = [bone_off[i], bone_off[i+1])
	const uint32_t *mtx_pos;   // per bone: position of the bone's transform
	const uint32_t *bind_idx;  // per bone: index into bind_pose
	const float *bind_pose;    // 12 floats each
	const uint32_t *bounds_idx; // per list
	const float *bounds6;      // interleaved MinMax table
	const float *world;
	const float4 *planes;      // 6 conditioned planes
	uint8_t *vis;              // per list: 1 = survives
	uint32_t n_lists;
};
__global__ void __launch_bounds__(256) k_cull_skinned(SkinCullDev d) {
	const uint32_t list = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
	if (list >= d.n_lists)
		return;
	const uint32_t b0 = d.bone_off[list], nb = min(d.bone_off[list + 1] - b0, 32u);
	float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f}, mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
	if (lane < nb) {
		const Mat4 m = mat4_mul(load_mat4(d.world + size_t(d.mtx_pos[b0 + lane]) * 12), load_mat4(d.bind_pose + size_t(d.bind_idx[b0 + lane]) * 12));
		const Aabb bb = minmax_transform(m, d.bounds6 + size_t(d.bounds_idx[list]) * 6);
#pragma unroll
		for (int k = 0; k < 3; ++k)
			mn[k] = bb.mn[k], mx[k] = bb.mx[k];
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1)
#pragma unroll
		for (int k = 0; k < 3; ++k) {
			mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], s));
			mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], s));
		}
	if (lane == 0) {
		Aabb u;
#pragma unroll
		for (int k = 0; k < 3; ++k)
			u.mn[k] = mn[k], u.mx[k] = mx[k];
		d.vis[list] = (nb == 0 || survives_frustum(d.planes, u)) ? 1 : 0; // a list without bones is never dropped
	}
}

// ---- "Enable" bool tracks ---------------------------------------------------------------------------------------------------------
// Scene::EvaluateBoundAnim, engine/scene.cpp:2158-2162: a bound "Enable" track is evaluated with EvaluateStep (engine/animation.h:130-139:
// the key at or before t, the first key before the track starts, the last one after it ends) and the node is enabled / disabled
// accordingly (EnableNode / DisableNode set or clear NF_Disabled, scene.cpp:472-503). On the device the node's state is the kEntDisabled
// bit of its tile meta row and of every display entry it owns; this kernel rewrites those words, before the level kernels of the frame
// read them. One thread per TARGET node: its tracks are walked in the reference's evaluation order, the last playing one wins.
struct EnableDev {
	const uint32_t *target_pos; // per target node
	const uint32_t *off;        // CSR: tracks of target i = [off[i], off[i+1])
	const uint4 *tracks;        // {key_off, key_count, play slot, -}
	const uint4 *keys;          // {t lo, t hi, value, -}
	const int64_t *play_t;
	uint32_t *tiles;
	uint32_t *ent_meta0, *ent_meta1; // per entry meta of the two tables (stand-alone cull, sort keys)
	const uint32_t *ent_off;         // overflow CSR by position
	uint2 *pent;
	uint32_t n;
};

__global__ void __launch_bounds__(128) k_sample_enable(EnableDev d) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= d.n)
		return;
	bool have = false, enable = true;
	for (uint32_t k = d.off[i]; k < d.off[i + 1]; ++k) {
		const uint4 trk = d.tracks[k];
		const int64_t at = d.play_t[trk.z];
		if (at == INT64_MIN || trk.y == 0)
			continue; // not playing / Evaluate() returns false on an empty track
		uint32_t lo = 0, hi = trk.y; // GetIntervalKeys: first key with key.t > at
		while (lo < hi) {
			const uint32_t mid = (lo + hi) >> 1;
			const uint4 key = d.keys[trk.x + mid];
			if (int64_t(uint64_t(key.x) | (uint64_t(key.y) << 32)) > at)
				hi = mid;
			else
				lo = mid + 1;
		}
		enable = d.keys[trk.x + (lo > 0 ? lo - 1 : 0)].z != 0;
		have = true;
	}
	if (!have)
		return;
	const uint32_t p = d.target_pos[i];
	uint32_t *row = d.tiles + tile_word(p, kRowEntMeta);
	const uint32_t meta = *row;
	*row = enable ? (meta & ~kEntDisabled) : (meta | kEntDisabled);
	if (meta & kEntHas) {
		uint32_t *m = ((meta & kEntKindTransparent) ? d.ent_meta1 : d.ent_meta0) + d.tiles[tile_word(p, kRowEntIdx)];
		*m = enable ? (*m & ~kEntDisabled) : (*m | kEntDisabled);
		if (meta & kEntMore)
			for (uint32_t k = d.ent_off[p]; k < d.ent_off[p + 1]; ++k) {
				uint2 rec = d.pent[k];
				rec.y = enable ? (rec.y & ~kEntDisabled) : (rec.y | kEntDisabled);
				d.pent[k] = rec;
				uint32_t *m2 = ((rec.x >> 31) ? d.ent_meta1 : d.ent_meta0) + (rec.x & 0x7fffffffu);
				*m2 = enable ? (*m2 & ~kEntDisabled) : (*m2 | kEntDisabled);
			}
	}
}

// disabled bit of the given positions (state read-back of the enable-tracked nodes)
__global__ void __launch_bounds__(128) k_gather_disabled(const uint32_t *__restrict__ tiles, const uint32_t *__restrict__ pos, uint32_t n, uint8_t *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n)
		out[i] = (tiles[tile_word(pos[i], kRowEntMeta)] & kEntDisabled) ? 1 : 0;
}

// ---- scene / instance bounding boxes ---------------------------------------------------------------------------------------------
// Scene::GetMinMax (engine/scene.cpp:1982) -> GetNodesMinMax (engine/node.cpp:265-285): union over the nodes that carry an Object of
// world[node transform] * Union(model.bounds[*]) (Node::GetMinMax node.cpp:106-119, Scene::GetObjectMinMax component.cpp:778-797,
// operator*(Mat4, MinMax) foundation/minmax.cpp:226-244, Union minmax.h). min/max are exact and order-independent, so the device
// reduces in POSITION order (coalesced world rows) with segment ids = scene instances; a warp whose 32 objects share a segment
// reduces by shuffles and issues 6 atomics, otherwise lanes are grouped with match.any. Floats are reduced as order-preserving
// unsigned keys (atomicMin/atomicMax on u32).
struct MinMaxDev {
	const uint32_t *obj_pos;   // per object: position of the node's transform, kRoot = no transform (identity world), ascending
	const uint32_t *obj_meta;  // per object: object-bounds index (bits 0..23)
	const uint32_t *obj_seg;   // per object: segment (instance) id, or null = one segment
	const float *world;
	const float *obounds6;     // interleaved {mn.x, mx.x, mn.y, mx.y, mn.z, mx.z} per model: union of its lists' bounds
	uint32_t *acc;             // per segment: 6 keys {mn.xyz (atomicMin), mx.xyz (atomicMax)} + count at [6]; 8 words per segment
	uint32_t n;
};

__device__ __forceinline__ uint32_t float_key(float f) {
	const uint32_t b = __float_as_uint(f);
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float key_float(uint32_t k) {
	const uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
	return __uint_as_float(b);
#else
	float f;
	memcpy(&f, &b, 4);
	return f;
#endif
}

__global__ void __launch_bounds__(256) k_minmax_init(uint32_t *acc, uint32_t n_seg) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n_seg * 8u) {
		const uint32_t w = i & 7u;
		acc[i] = w < 3 ? 0xffffffffu : 0u;
	}
}

// Persistent grid (a few CTAs per SM): every WARP owns a contiguous chunk of the position-sorted objects and walks it 32 at a time
// (32 world rows = 1536 contiguous bytes when every node carries an object). Objects of one segment form long runs in position order
// (a level of one instance), so a warp folds each uniform group of 32 with a shuffle butterfly into a running box and touches memory
// only when the segment changes; the 8 warps of a CTA merge their last runs in shared memory. One-segment scenes end with 7 atomics
// per CTA; mixed warps (segment boundaries) publish their groups directly.
__device__ __forceinline__ void minmax_publish(uint32_t *acc, uint32_t seg, const uint32_t (&k)[6], uint32_t cnt) {
	uint32_t *a = acc + size_t(seg) * 8;
#pragma unroll
	for (int j = 0; j < 3; ++j) {
		atomicMin(a + j, k[j]);
		atomicMax(a + 3 + j, k[3 + j]);
	}
	atomicAdd(a + 6, cnt);
}

__global__ void __launch_bounds__(256) k_minmax(MinMaxDev d, uint32_t per_warp) {
	__shared__ uint32_t s_run[8][8]; // per warp: seg, 6 keys, count
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t gw = blockIdx.x * 8u + warp;
	const uint64_t b0 = uint64_t(gw) * per_warp;
	const uint32_t begin = uint32_t(min(b0, uint64_t(d.n))), end = uint32_t(min(b0 + per_warp, uint64_t(d.n)));
	uint32_t run_seg = 0xffffffffu, run_cnt = 0;
	uint32_t run[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
	for (uint32_t base = begin; base < end; base += 32) {
		const uint32_t i = base + lane;
		const bool live = i < end;
		uint32_t seg = 0xffffffffu;
		uint32_t k[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
		if (live) {
			seg = d.obj_seg ? d.obj_seg[i] : 0u;
			const uint32_t p = d.obj_pos[i];
			Mat4 w;
			if (p != kRoot) {
				w = load_mat4(d.world + size_t(p) * 12);
			} else { // Transform::GetWorld on an invalid transform returns Mat4::Identity (component.cpp:250-256)
#pragma unroll
				for (int r = 0; r < 3; ++r)
#pragma unroll
					for (int c = 0; c < 4; ++c)
						w.m[r][c] = r == c ? 1.f : 0.f;
			}
			const Aabb bb = minmax_transform(w, d.obounds6 + size_t(d.obj_meta[i] & 0xffffffu) * 6);
#pragma unroll
			for (int j = 0; j < 3; ++j)
				k[j] = float_key(bb.mn[j]), k[3 + j] = float_key(bb.mx[j]);
		}
		// uniform when every LIVE lane has the segment of the first lane (dead lanes carry neutral keys)
		const uint32_t seg0 = __shfl_sync(0xffffffffu, seg, 0);
		if (__all_sync(0xffffffffu, !live || seg == seg0)) {
			uint32_t cnt = live ? 1u : 0u;
#pragma unroll
			for (int s = 16; s > 0; s >>= 1) {
#pragma unroll
				for (int j = 0; j < 3; ++j) {
					k[j] = min(k[j], __shfl_xor_sync(0xffffffffu, k[j], s));
					k[3 + j] = max(k[3 + j], __shfl_xor_sync(0xffffffffu, k[3 + j], s));
				}
				cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
			}
			if (seg0 != run_seg) { // a new run: the finished one goes to memory
				if (run_cnt && lane == 0)
					minmax_publish(d.acc, run_seg, run, run_cnt);
				run_seg = seg0, run_cnt = 0;
#pragma unroll
				for (int j = 0; j < 3; ++j)
					run[j] = 0xffffffffu, run[3 + j] = 0u;
			}
#pragma unroll
			for (int j = 0; j < 3; ++j)
				run[j] = min(run[j], k[j]), run[3 + j] = max(run[3 + j], k[3 + j]);
			run_cnt += cnt;
		} else { // a segment boundary inside the 32: every group folds into its lowest lane, which publishes
			const uint32_t peers = __match_any_sync(0xffffffffu, seg);
			const uint32_t leader = __ffs(peers) - 1;
			uint32_t rest = peers & ~(1u << leader);
			uint32_t cnt = live ? 1u : 0u;
			const uint32_t trips = __reduce_max_sync(0xffffffffu, uint32_t(__popc(rest)));
			for (uint32_t t = 0; t < trips; ++t) { // all lanes take part in every shuffle; a lane whose list is exhausted reads itself
				const uint32_t src = rest ? __ffs(rest) - 1 : lane;
				rest &= rest - 1;
#pragma unroll
				for (int j = 0; j < 3; ++j) {
					k[j] = min(k[j], __shfl_sync(0xffffffffu, k[j], src));
					k[3 + j] = max(k[3 + j], __shfl_sync(0xffffffffu, k[3 + j], src));
				}
				const uint32_t c2 = __shfl_sync(0xffffffffu, live ? 1u : 0u, src);
				if (src != lane)
					cnt += c2;
			}
			if (lane == leader && live)
				minmax_publish(d.acc, seg, k, cnt);
		}
	}
	// the last run of every warp: merged across the CTA when neighbours share the segment
	if (lane == 0) {
		s_run[warp][0] = run_cnt ? run_seg : 0xffffffffu;
#pragma unroll
		for (int j = 0; j < 6; ++j)
			s_run[warp][1 + j] = run[j];
		s_run[warp][7] = run_cnt;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t seg = 0xffffffffu, cnt = 0;
		uint32_t k[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
		for (int w = 0; w < 8; ++w) {
			const uint32_t sw = s_run[w][0];
			if (sw == 0xffffffffu)
				continue;
			if (sw != seg) {
				if (cnt)
					minmax_publish(d.acc, seg, k, cnt);
				seg = sw, cnt = 0;
#pragma unroll
				for (int j = 0; j < 3; ++j)
					k[j] = 0xffffffffu, k[3 + j] = 0u;
			}
#pragma unroll
			for (int j = 0; j < 3; ++j)
				k[j] = min(k[j], s_run[w][1 + j]), k[3 + j] = max(k[3 + j], s_run[w][4 + j]);
			cnt += s_run[w][7];
		}
		if (cnt)
			minmax_publish(d.acc, seg, k, cnt);
	}
}

// ---- transparent-list sort keys -------------------------------------------------------------------------------------------
// ComputeModelDisplayListSortKeys (engine/scene_forward_pipeline.cpp:438-483) for the SURVIVORS of a culled list: model-view matrix =
// view * world (operator*(Mat4,Mat4)), the 8 corners of the list's MinMax through operator*(Mat4,Vec3) (foundation/matrix4.cpp:78-81;
// only z is used), the closest z clamped at 0, then ComputeSortKey (engine/render_pipeline.cpp:89-94): +100 m, clamp, 1 mm fixed point.
struct SortKeyDev {
	const uint32_t *visible;  // surviving entry indices
	const uint32_t *count;    // how many (device word)
	const uint32_t *mtx_pos;  // per entry
	const uint32_t *meta;     // per entry: bounds idx | flags
	const float *world;
	const float *bounds6;     // interleaved {mn.x, mx.x, mn.y, mx.y, mn.z, mx.z}
	float view[12];           // ViewState::view
	uint32_t *keys;           // out: one per survivor
};

__global__ void __launch_bounds__(256) k_sort_keys(SortKeyDev d, uint32_t capacity) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= capacity || i >= *d.count)
		return;
	const uint32_t e = d.visible[i];
	Mat4 view;
#pragma unroll
	for (int k = 0; k < 12; ++k)
		view.m[k / 4][k % 4] = d.view[k];
	const Mat4 mv = mat4_mul(view, load_mat4(d.world + size_t(d.mtx_pos[e]) * 12));
	const float *b = d.bounds6 + size_t(d.meta[e] & 0xffffffu) * 6;
	const float mnx = b[0], mxx = b[1], mny = b[2], mxy = b[3], mnz = b[4], mxz = b[5];
	// corner order of the reference (the running minimum is order-independent for finite values; kept anyway)
	const float cx[8] = {mnx, mxx, mxx, mnx, mnx, mxx, mxx, mnx};
	const float cy[8] = {mny, mny, mxy, mxy, mny, mny, mxy, mxy};
	const float cz[8] = {mnz, mnz, mnz, mnz, mxz, mxz, mxz, mxz};
	float closest = 3.402823466e+38f;
#pragma unroll
	for (int k = 0; k < 8; ++k) {
		const float z = cx[k] * mv.m[2][0] + cy[k] * mv.m[2][1] + cz[k] * mv.m[2][2] + mv.m[2][3];
		if (z < closest)
			closest = z;
	}
	if (closest < 0.f)
		closest = 0.f;
	float depth = closest + 100.f;
	if (depth < 0.f)
		depth = 0.f;
	d.keys[i] = uint32_t(__float2ll_rz(depth * 1000.f)); // x86-64 converts through a 64-bit integer and keeps the low word
}

// ---- Scene::SetTransformWorldMatrix / SetTransformLocalMatrix, batched --------------------------------------------------------------
// engine/component.cpp:265-281 / :236-248: the world row is written, local = InverseFast(world[parent]) * world (or the given local,
// with world = world[parent] * local) and Decompose(local) (foundation/matrix4.cpp:208-252) refreshes the transform's TRS, Euler angles
// through ToEuler. The host splits a batch into launches such that no entry reads or rewrites a row another entry of the same launch
// writes, so a batch behaves like the reference's sequential calls.
__device__ __forceinline__ Mat4 inverse_fast(const Mat4 &m) { // foundation/matrix4.cpp:293-303
	float sc[3];
#pragma unroll
	for (int c = 0; c < 3; ++c)
		sc[c] = 1.F / (m.m[0][c] * m.m[0][c] + m.m[1][c] * m.m[1][c] + m.m[2][c] * m.m[2][c]);
	Mat4 o;
#pragma unroll
	for (int r = 0; r < 3; ++r)
#pragma unroll
		for (int c = 0; c < 3; ++c)
			o.m[r][c] = m.m[c][r] * sc[r];
	const float tx = -m.m[0][3], ty = -m.m[1][3], tz = -m.m[2][3];
#pragma unroll
	for (int r = 0; r < 3; ++r)
		o.m[r][3] = o.m[r][0] * tx + o.m[r][1] * ty + o.m[r][2] * tz;
	return o;
}

__global__ void __launch_bounds__(128) k_set_matrices(const uint32_t *__restrict__ pos, const float4 *__restrict__ mats, uint32_t n, int is_local,
	float *__restrict__ world, uint32_t *__restrict__ tiles, float4 *__restrict__ anim_val) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const uint32_t p = pos[i];
	const float4 r0 = mats[i * 3], r1 = mats[i * 3 + 1], r2 = mats[i * 3 + 2];
	Mat4 given;
	given.m[0][0] = r0.x, given.m[0][1] = r0.y, given.m[0][2] = r0.z, given.m[0][3] = r0.w;
	given.m[1][0] = r1.x, given.m[1][1] = r1.y, given.m[1][2] = r1.z, given.m[1][3] = r1.w;
	given.m[2][0] = r2.x, given.m[2][1] = r2.y, given.m[2][2] = r2.z, given.m[2][3] = r2.w;
	uint32_t *d = tiles + tile_word(p, 0);
	const uint32_t par = d[kRowParent * 32];
	Mat4 local = given, wrld = given;
	if (par != kRoot) {
		const Mat4 pw = load_mat4(world + size_t(par) * 12);
		if (is_local)
			wrld = mat4_mul(pw, given);
		else
			local = mat4_mul(inverse_fast(pw), given);
	}
	store_mat4(world + size_t(p) * 12, wrld);
	// Decompose(local): position, scale (negative scale folded into x), renormalised axes -> Euler YXZ
	const float vx[3] = {local.m[0][0], local.m[1][0], local.m[2][0]}, vy[3] = {local.m[0][1], local.m[1][1], local.m[2][1]},
				vz[3] = {local.m[0][2], local.m[1][2], local.m[2][2]};
	float sx = sqrtf(vx[0] * vx[0] + vx[1] * vx[1] + vx[2] * vx[2]);
	const float sy = sqrtf(vy[0] * vy[0] + vy[1] * vy[1] + vy[2] * vy[2]);
	const float sz = sqrtf(vz[0] * vz[0] + vz[1] * vz[1] + vz[2] * vz[2]);
	const float lx = vy[1] * vz[2] - vy[2] * vz[1], ly = vy[2] * vz[0] - vy[0] * vz[2], lz = vy[0] * vz[1] - vy[1] * vz[0]; // Cross(vy, vz)
	if (lx * vx[0] + ly * vx[1] + lz * vx[2] < 0.f)
		sx = -sx;
	float m[3][3]; // column c = axis c / its scale (identity axis when the scale is 0)
#pragma unroll
	for (int r = 0; r < 3; ++r) {
		m[r][0] = sx ? vx[r] / sx : (r == 0 ? 1.f : 0.f);
		m[r][1] = sy ? vy[r] / sy : (r == 1 ? 1.f : 0.f);
		m[r][2] = sz ? vz[r] / sz : (r == 2 ? 1.f : 0.f);
	}
	float ex, ey, ez;
	mat3_to_euler_yxz(m[0][0], m[0][2], m[1][0], m[1][1], m[1][2], m[2][0], m[2][2], ex, ey, ez);
	const float trs[9] = {local.m[0][3], local.m[1][3], local.m[2][3], ex, ey, ez, sx, sy, sz};
	const uint32_t meta = d[kRowEntMeta * 32];
#pragma unroll
	for (int g = 0; g < 3; ++g) {
		if (meta & (kLaneAnimated << g)) // track-driven triple: the value lives in anim_val
			anim_val[d[g * 96]] = make_float4(trs[g * 3], trs[g * 3 + 1], trs[g * 3 + 2], 0.f);
		else
			d[g * 96] = __float_as_uint(trs[g * 3]), d[g * 96 + 32] = __float_as_uint(trs[g * 3 + 1]), d[g * 96 + 64] = __float_as_uint(trs[g * 3 + 2]);
	}
}

// ---- layout shuffles -----------------------------------------------------------------------------------------------------
// host TRS upload: AoS hg::TransformTRS (9 floats, engine/node.h:36-38) in SLOT order -> tile rows in POSITION order
__global__ void __launch_bounds__(256) k_scatter_trs(const float *__restrict__ staging9, const uint32_t *__restrict__ pos_of_slot, uint32_t first_slot,
	uint32_t count, uint32_t *__restrict__ tiles, float4 *__restrict__ anim_val) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	const uint32_t p = pos_of_slot ? pos_of_slot[first_slot + i] : first_slot + i;
	const float *s = staging9 + size_t(i) * 9;
	uint32_t *d = tiles + tile_word(p, 0);
	const uint32_t meta = d[kRowEntMeta * 32];
#pragma unroll
	for (int g = 0; g < 3; ++g) {
		if (meta & (kLaneAnimated << g)) // track-driven triple: the value lives in anim_val (SetTransformPos on an animated node)
			anim_val[d[g * 96]] = make_float4(s[g * 3], s[g * 3 + 1], s[g * 3 + 2], 0.f);
		else
			d[g * 96] = __float_as_uint(s[g * 3]), d[g * 96 + 32] = __float_as_uint(s[g * 3 + 1]), d[g * 96 + 64] = __float_as_uint(s[g * 3 + 2]);
	}
}

__global__ void __launch_bounds__(256) k_gather_trs(const uint32_t *__restrict__ tiles, const uint32_t *__restrict__ pos_of_slot, uint32_t first_slot,
	uint32_t count, float *__restrict__ out9, const float4 *__restrict__ anim_val) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	const uint32_t p = pos_of_slot ? pos_of_slot[first_slot + i] : first_slot + i;
	const uint32_t *s = tiles + tile_word(p, 0);
	const uint32_t meta = s[kRowEntMeta * 32];
	float *o = out9 + size_t(i) * 9;
#pragma unroll
	for (int g = 0; g < 3; ++g) {
		if (meta & (kLaneAnimated << g)) {
			const float4 v = anim_val[s[g * 96]];
			o[g * 3] = v.x, o[g * 3 + 1] = v.y, o[g * 3 + 2] = v.z;
		} else {
			o[g * 3] = __uint_as_float(s[g * 96]), o[g * 3 + 1] = __uint_as_float(s[g * 96 + 32]), o[g * 3 + 2] = __uint_as_float(s[g * 96 + 64]);
		}
	}
}

// Bind / unbind the track-driven triples. bind: anim_val[slot] takes the triple's current value, the tile's x row takes the slot
// and the animated bit is set; unbind is the inverse (the last sampled value stays in the TRS, as in the reference where
// EvaluateBoundAnim writes into the Transform, engine/scene.cpp:2165-2180).
__global__ void __launch_bounds__(256) k_bind_lanes(uint32_t *__restrict__ tiles, float4 *__restrict__ anim_val, const uint32_t *__restrict__ bind_pos,
	const uint8_t *__restrict__ bind_group, uint32_t n, int bind) {
	const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
	if (slot >= n)
		return;
	const uint32_t g = bind_group[slot];
	uint32_t *d = tiles + tile_word(bind_pos[slot], 0);
	if (bind) {
		anim_val[slot] = make_float4(__uint_as_float(d[g * 96]), __uint_as_float(d[g * 96 + 32]), __uint_as_float(d[g * 96 + 64]), 0.f);
		d[g * 96] = slot;
		atomicOr(d + kRowEntMeta * 32, kLaneAnimated << g);
	} else {
		const float4 v = anim_val[slot];
		d[g * 96] = __float_as_uint(v.x), d[g * 96 + 32] = __float_as_uint(v.y), d[g * 96 + 64] = __float_as_uint(v.z);
		atomicAnd(d + kRowEntMeta * 32, ~(kLaneAnimated << g));
	}
}

// rows 9..11 of the tiles (parent position, first entry meta, first entry index) from position-ordered staging arrays
__global__ void __launch_bounds__(256) k_fill_rows(uint32_t *__restrict__ tiles, uint32_t row, const uint32_t *__restrict__ src, uint32_t n_pos) {
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p < n_pos)
		tiles[tile_word(p, row)] = src[p];
}

// ---- incremental edits (hgb_append_nodes / hgb_remove_nodes / hgb_append_display_entries / hgb_set_entry_flags) --------------------
// A few threads each: the point is that nothing else moves — no re-sort, no table rebuild, no graph capture.
struct NodeWrite {
	uint32_t pos, parent; // parent position, kRoot or kDead (free the position)
	float trs[9];
	uint32_t zero_world;  // 1: new node — both world rows start from zero like a std::vector::resize'd slot
};
__global__ void __launch_bounds__(128) k_write_nodes(uint32_t *__restrict__ tiles, float *__restrict__ world0, float *__restrict__ world1, uint8_t *__restrict__ preset,
	const NodeWrite *__restrict__ w, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const NodeWrite x = w[i];
	uint32_t *t = tiles + tile_word(x.pos, 0);
#pragma unroll
	for (int r = 0; r < 9; ++r)
		t[r * 32] = __float_as_uint(x.trs[r]);
	t[kRowParent * 32] = x.parent;
	t[kRowEntMeta * 32] = 0;
	t[kRowEntIdx * 32] = 0;
	preset[x.pos] = 0;
	if (x.zero_world)
		for (int k = 0; k < 12; ++k)
			world0[size_t(x.pos) * 12 + k] = 0.f, world1[size_t(x.pos) * 12 + k] = 0.f;
}

// A removed node: its display entries are disabled for good (entry tables, tile row, overflow records; visibility bytes of both
// parities cleared, since nobody rewrites them once the node is gone). In the SORTED region the position stays until the next depth
// sort as a childless root that owns a disabled entry; in the append region it becomes a free position (kDead).
struct RetireDev {
	uint32_t *tiles;
	uint32_t *ent_meta[2];
	uint8_t *vis[2][2]; // [kind][parity]
	const uint32_t *ent_off;
	uint2 *pent;
	const uint32_t *pos;
	uint32_t n, dead_from;
};
__global__ void __launch_bounds__(128) k_retire_nodes(RetireDev d) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= d.n)
		return;
	const uint32_t p = d.pos[i];
	uint32_t *row = d.tiles + tile_word(p, kRowEntMeta);
	const uint32_t m = *row;
	if ((m & kEntHas) && d.ent_meta[0]) {
		const int k = (m & kEntKindTransparent) ? 1 : 0;
		const uint32_t e = d.tiles[tile_word(p, kRowEntIdx)];
		d.ent_meta[k][e] |= kEntDisabled;
		d.vis[k][0][e] = 0, d.vis[k][1][e] = 0;
		if (m & kEntMore)
			for (uint32_t j = d.ent_off[p]; j < d.ent_off[p + 1]; ++j) {
				const uint2 rec = d.pent[j];
				const int k2 = int(rec.x >> 31);
				d.pent[j].y = rec.y | kEntDisabled;
				d.ent_meta[k2][rec.x & 0x7fffffffu] |= kEntDisabled;
				d.vis[k2][0][rec.x & 0x7fffffffu] = 0, d.vis[k2][1][rec.x & 0x7fffffffu] = 0;
			}
	}
	if (p >= d.dead_from) {
		d.tiles[tile_word(p, kRowParent)] = kDead;
		*row = 0;
	} else {
		d.tiles[tile_word(p, kRowParent)] = kRoot;
		*row = m | kEntDisabled; // bound triples keep their anim_val slots (the sampler still writes them)
	}
}

struct EntryWrite {
	uint32_t entry, pos, meta; // meta: bounds index | kEntDisabled
};
// new display entries at the end of a table, each the FIRST entry of its node (rows 10 / 11 of the node's tile)
__global__ void __launch_bounds__(128) k_append_entries(uint32_t *__restrict__ tiles, uint32_t *__restrict__ mtx_pos, uint32_t *__restrict__ ent_meta, uint8_t *__restrict__ vis_a,
	uint8_t *__restrict__ vis_b, const EntryWrite *__restrict__ w, uint32_t n, uint32_t kind_bit) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const EntryWrite x = w[i];
	mtx_pos[x.entry] = x.pos;
	ent_meta[x.entry] = x.meta;
	vis_a[x.entry] = 0, vis_b[x.entry] = 0;
	uint32_t *row = tiles + tile_word(x.pos, kRowEntMeta);
	*row = (*row & kLaneAnimatedMask) | x.meta | kEntHas | kind_bit;
	tiles[tile_word(x.pos, kRowEntIdx)] = x.entry;
}

// disabled flag of entries [first, first + n) of one table: the entry table (stand-alone cull) and wherever the fused pass reads it —
// the node's tile row when it is the node's first entry, else its record in the overflow CSR
__global__ void __launch_bounds__(128) k_set_entry_flags(uint32_t *__restrict__ tiles, const uint32_t *__restrict__ mtx_pos, uint32_t *__restrict__ ent_meta,
	const uint32_t *__restrict__ ent_off, uint2 *__restrict__ pent, const uint8_t *__restrict__ disabled, uint32_t first, uint32_t n, uint32_t kind_bit) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const uint32_t e = first + i, dis = disabled[i] ? kEntDisabled : 0u;
	ent_meta[e] = (ent_meta[e] & ~kEntDisabled) | dis;
	const uint32_t p = mtx_pos[e];
	uint32_t *row = tiles + tile_word(p, kRowEntMeta);
	const uint32_t m = *row;
	if ((m & kEntHas) && (m & kEntKindTransparent) == kind_bit && tiles[tile_word(p, kRowEntIdx)] == e) {
		*row = (m & ~kEntDisabled) | dis;
	} else if (m & kEntMore) {
		for (uint32_t k = ent_off[p]; k < ent_off[p + 1]; ++k)
			if (pent[k].x == (e | kind_bit))
				pent[k].y = (pent[k].y & ~kEntDisabled) | dis;
	}
}

// visible entry indices -> the caller's 16-byte records of those entries (hgb_get_visible_records_ptr): the sparse gather runs here,
// where random 16-byte reads are cheap, and the host receives one contiguous array
__global__ void __launch_bounds__(256) k_gather_records(const uint4 *__restrict__ table, const uint32_t *__restrict__ visible, const uint32_t *__restrict__ count,
	uint32_t capacity, uint4 *__restrict__ out) {
	const uint32_t n = min(*count, capacity);
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
		out[i] = table[visible[i]];
}

// re-parenting inside a level (hgb_set_parents): only the parent row of the moved positions changes
__global__ void __launch_bounds__(128) k_set_parent_rows(uint32_t *__restrict__ tiles, const uint32_t *__restrict__ pos, const uint32_t *__restrict__ par, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n)
		tiles[tile_word(pos[i], kRowParent)] = par[i];
}

// worlds in POSITION order -> SLOT order (only when the permutation is not the identity); one thread per float4
__global__ void __launch_bounds__(256) k_gather_worlds(const float4 *__restrict__ world, const uint32_t *__restrict__ pos_of_slot, uint32_t first_slot,
	uint32_t count, float4 *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count * 3)
		return;
	const uint32_t s = i / 3, r = i - s * 3;
	out[i] = world[size_t(pos_of_slot ? pos_of_slot[first_slot + s] : first_slot + s) * 3 + r];
}

// the inverse: SLOT order -> POSITION order (hierarchy edits carry both world buffers over to the new order)
__global__ void __launch_bounds__(256) k_scatter_worlds(const float4 *__restrict__ in, const uint32_t *__restrict__ pos_of_slot, uint32_t count,
	float4 *__restrict__ world) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count * 3)
		return;
	const uint32_t s = i / 3, r = i - s * 3;
	world[size_t(pos_of_slot ? pos_of_slot[s] : s) * 3 + r] = in[i];
}

// SetNodeWorldMatrix overrides (engine/scene.cpp:768-786): write the matrix and mark the position as updated for this frame
__global__ void k_apply_presets(const uint32_t *__restrict__ pos, const float4 *__restrict__ mats, uint32_t n, float4 *__restrict__ world,
	uint8_t *__restrict__ preset) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const uint32_t p = pos[i];
	world[size_t(p) * 3 + 0] = mats[i * 3 + 0];
	world[size_t(p) * 3 + 1] = mats[i * 3 + 1];
	world[size_t(p) * 3 + 2] = mats[i * 3 + 2];
	preset[p] = 1;
}

} // namespace hgb
