// Animation key sampling, skin bone matrices and the layout shuffles around host uploads/downloads.
#pragma once
#include "hgb_kernels.cuh"

namespace hgb {

// ---- animation key sampling -----------------------------------------------------------------------------------------
// One thread per bound track. Replaces Scene::EvaluateBoundAnim(BoundToNodeAnim) (engine/scene.cpp:2154-2181) for the
// transform targets: Evaluate(track, t, value) -> TRS lane of the bound node's transform.
//   GetIntervalKeys (engine/animation.h:109-128) is a linear scan for the first key with key.t > t; the binary search
//   below returns the same index (upper_bound) because keys are sorted by time.
struct Track { // 16 B
	uint32_t target_pos; // position of the target transform
	uint32_t key_off;    // first key in the pool of its kind
	uint32_t key_count;
	uint16_t play;       // index into play_t[]
	uint8_t kind;        // 0 Position, 1 Rotation (Euler vec3), 2 Scale, 3 Rotation (quaternion)
	uint8_t pad;
};

struct AnimDev {
	const Track *tracks;
	const int64_t *play_t;    // per playing anim: evaluation time (ns), INT64_MIN = not playing
	const int64_t *v3_t;      // vec3 Hermite key pool: times
	const float4 *v3_val;     //   {v.x, v.y, v.z, tension}
	const float *v3_bias;     //   bias
	const int64_t *q_t;       // quaternion key pool
	const float4 *q_val;
	float *trs;               // SoA TRS (see SceneDev)
	size_t trs_stride;
};

__device__ __forceinline__ float time_to_sec_f(int64_t t) { return __double2float_rn(__ddiv_rn(double(t), 1000000000.0)); } // foundation/time.h:13

__device__ __forceinline__ uint32_t upper_bound_i64(const int64_t *__restrict__ t, uint32_t n, int64_t at) {
	uint32_t lo = 0, hi = n;
	while (lo < hi) {
		const uint32_t mid = (lo + hi) >> 1;
		if (t[mid] > at)
			hi = mid;
		else
			lo = mid + 1;
	}
	return lo;
}

// HermiteInterpolate foundation/math.h:137-149 on one component
__device__ __forceinline__ float hermite1(float y0, float y1, float y2, float y3, float a0, float a1, float a2, float a3, float tension, float bias) {
	float m0 = (y1 - y0) * (1.F + bias) * (1.F - tension) / 2.F;
	m0 += (y2 - y1) * (1.F - bias) * (1.F - tension) / 2.F;
	float m1 = (y2 - y1) * (1.F + bias) * (1.F - tension) / 2.F;
	m1 += (y3 - y2) * (1.F - bias) * (1.F - tension) / 2.F;
	return y1 * a0 + m0 * a1 + m1 * a2 + y2 * a3;
}

// ToEuler(Normalize(q)) engine/scene.cpp:2172 : foundation/quaternion.cpp:57-60 (Normalize), :141-166 (ToMatrix3),
// foundation/matrix3.cpp:91-137 (ToEuler, euler_order_yxz: i=1 j=0 k=2). atan2 is the double libm function in the reference
// (float args promoted, result narrowed); CUDA's double atan2 (<= 2 ulp) narrows to the same float except when the double
// result sits within ~1e-16 relative of a float rounding boundary — covered by the 1e-5 matrix tolerance.
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
	const float s1 = -m12;
	const float c1 = sqrtf(m11 * m11 + m10 * m10);
	float a0, a2;
	const float a1 = __double2float_rn(atan2(double(s1), double(c1)));
	if (c1 > 1.1920929e-07f) {
		a0 = __double2float_rn(atan2(double(m02), double(m22)));
		a2 = __double2float_rn(atan2(double(m10), double(m11)));
	} else {
		a0 = 0.F;
		const float sign_s1 = s1 < 0.F ? -1.F : (s1 > 0.F ? 1.F : 0.F);
		a2 = __double2float_rn(__dmul_rn(double(-sign_s1), atan2(double(-m20), double(m00))));
	}
	ex = a1, ey = a0, ez = a2;
}

__global__ void __launch_bounds__(256) k_sample_tracks(AnimDev a, uint32_t first, uint32_t n_tracks) {
	const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= first + n_tracks)
		return;
	const Track trk = a.tracks[i];
	const int64_t at = a.play_t[trk.play];
	if (at == INT64_MIN || trk.key_count == 0) // not playing / Evaluate() returns false on an empty track (animation.h:156)
		return;
	const uint32_t n = trk.key_count;
	float *dst = a.trs + trk.target_pos;
	if (trk.kind == 3) { // AnimTrackT<Quaternion>: EvaluateLinear animation.h:141-153, then ToEuler(Normalize())
		const int64_t *t = a.q_t + trk.key_off;
		const float4 *v = a.q_val + trk.key_off;
		const uint32_t ub = upper_bound_i64(t, n, at);
		float4 q;
		if (ub == 0 || ub == n) {
			q = v[ub == 0 ? 0 : n - 1];
		} else {
			const float4 q0 = v[ub - 1], q1 = v[ub];
			const float k = time_to_sec_f(at - t[ub - 1]) / time_to_sec_f(t[ub] - t[ub - 1]);
			q = make_float4(q0.x + (q1.x - q0.x) * k, q0.y + (q1.y - q0.y) * k, q0.z + (q1.z - q0.z) * k, q0.w + (q1.w - q0.w) * k);
		}
		float ex, ey, ez;
		quat_to_euler(q.x, q.y, q.z, q.w, ex, ey, ez);
		dst[3 * a.trs_stride] = ex, dst[4 * a.trs_stride] = ey, dst[5 * a.trs_stride] = ez;
		return;
	}
	// AnimTrackHermiteT<Vec3>: EvaluateHermite animation.h:155-169
	const int64_t *t = a.v3_t + trk.key_off;
	const float4 *v = a.v3_val + trk.key_off;
	const uint32_t ub = upper_bound_i64(t, n, at);
	float ox, oy, oz;
	if (ub == 0 || ub == n) {
		const float4 k = v[ub == 0 ? 0 : n - 1];
		ox = k.x, oy = k.y, oz = k.z;
	} else {
		const uint32_t kf1 = ub - 1, kf2 = ub;
		const uint32_t kf0 = kf1 > 0 ? kf1 - 1 : 0, kf3 = kf2 + 1 < n - 1 ? kf2 + 1 : n - 1;
		const float u = time_to_sec_f(at - t[kf1]) / time_to_sec_f(t[kf2] - t[kf1]);
		const float4 y0 = v[kf0], y1 = v[kf1], y2 = v[kf2], y3 = v[kf3];
		const float tension = y1.w, bias = a.v3_bias[trk.key_off + kf1];
		const float t2 = u * u, t3 = t2 * u;
		const float a0 = 2.F * t3 - 3.F * t2 + 1.F;
		const float a1 = t3 - 2.F * t2 + u;
		const float a2 = t3 - t2;
		const float a3 = -2.F * t3 + 3.F * t2;
		ox = hermite1(y0.x, y1.x, y2.x, y3.x, a0, a1, a2, a3, tension, bias);
		oy = hermite1(y0.y, y1.y, y2.y, y3.y, a0, a1, a2, a3, tension, bias);
		oz = hermite1(y0.z, y1.z, y2.z, y3.z, a0, a1, a2, a3, tension, bias);
	}
	const size_t lane0 = size_t(trk.kind) * 3;
	dst[lane0 * a.trs_stride] = ox, dst[(lane0 + 1) * a.trs_stride] = oy, dst[(lane0 + 2) * a.trs_stride] = oz;
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

// ---- layout shuffles -----------------------------------------------------------------------------------------------------
// host TRS upload: AoS hg::TransformTRS (9 floats, engine/node.h:36-38) in SLOT order -> SoA lanes in POSITION order
__global__ void __launch_bounds__(256) k_scatter_trs(const float *__restrict__ staging9, const uint32_t *__restrict__ pos_of_slot, uint32_t first_slot,
	uint32_t count, float *__restrict__ trs, size_t stride) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	const uint32_t p = pos_of_slot ? pos_of_slot[first_slot + i] : first_slot + i;
	const float *s = staging9 + size_t(i) * 9;
#pragma unroll
	for (int c = 0; c < 9; ++c)
		trs[c * stride + p] = s[c];
}

__global__ void __launch_bounds__(256) k_gather_trs(const float *__restrict__ trs, size_t stride, const uint32_t *__restrict__ pos_of_slot, uint32_t first_slot,
	uint32_t count, float *__restrict__ out9) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	const uint32_t p = pos_of_slot ? pos_of_slot[first_slot + i] : first_slot + i;
#pragma unroll
	for (int c = 0; c < 9; ++c)
		out9[size_t(i) * 9 + c] = trs[c * stride + p];
}

// worlds in POSITION order -> SLOT order (only when the permutation is not the identity); one thread per float4
__global__ void __launch_bounds__(256) k_gather_worlds(const float4 *__restrict__ world, const uint32_t *__restrict__ pos_of_slot, uint32_t first_slot,
	uint32_t count, float4 *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count * 3)
		return;
	const uint32_t s = i / 3, r = i - s * 3;
	out[i] = world[size_t(pos_of_slot[first_slot + s]) * 3 + r];
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
