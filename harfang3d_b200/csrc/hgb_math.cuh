// Device arithmetic of the scene-update path, written so that every result is bit-identical to the reference's
// g++ -O2 x86-64 build (SSE2 scalar float, no contraction) on the same inputs:
//   * this TU is compiled with -fmad=false, so no float/double mul+add is ever contracted; the only fused operations
//     are the explicit fma() calls below, which mirror glibc's -mfma build of sinf/cosf;
//   * division and sqrt are the IEEE round-to-nearest forms (nvcc defaults -prec-div=true -prec-sqrt=true, -ftz=false).
// Reference anchors are given per function (paths relative to the reference's harfang/ directory).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hgb {

struct Mat4 { // == hg::Mat4::m, row-major 3x4 (foundation/matrix4.h:25)
	float m[3][4];
};

// ---- packed FP32 (Blackwell mul.f32x2: two IEEE round-to-nearest products per issue slot) ----------------------------------------
// The level kernel is co-limited by instruction issue (profiles/README.md r01b), and a third of its instructions are FMULs that come
// in pairs sharing one operand. mul.rn.f32x2 computes two independent correctly-rounded products, so every result keeps the bits of
// the scalar form. ONLY products are packed, and packed sums are used only where no product feeds them directly: ptxas 12.9 contracts
// mul.rn.f32x2 -> add.rn.f32x2 chains into FFMA2 even under -fmad=false (verified in SASS), which would break bit-exactness.
// tests/test_capi_symbols.py asserts that the library's SASS holds no FFMA/FFMA2 at all.
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t pk2(float lo, float hi) {
	f32x2_t r;
	asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
	return r;
}
__device__ __forceinline__ void upk2(f32x2_t v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2_t mul2(f32x2_t a, f32x2_t b) {
	f32x2_t r;
	asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ f32x2_t add2(f32x2_t a, f32x2_t b) {
	f32x2_t r;
	asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}

// ---- sinf/cosf --------------------------------------------------------------------------------------------------
// The reference calls libm sinf/cosf (foundation/matrix3.cpp:199-204). glibc 2.39 (sysdeps/ieee754/flt-32/sincosf.h)
// reduces in double, evaluates a double polynomial and rounds once to float; on FMA-capable hosts each a*b+c is one
// fused op. Restating that in FP64 here gives the same float for all 2^32 inputs (host-side proof: oracle/exhaustive_sincos.c),
// which is what makes visibility bit-exact instead of "1 ulp away near a frustum plane". B200 has a full-rate FP64 pipe;
// ~16 DP ops per angle stay hidden under the HBM traffic of the kernel.
__device__ __forceinline__ uint32_t abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ffu; }

// 4/pi as a sliding window of 32-bit words, for |x| >= 120 (rare: Euler angles are normally within a few turns)
__device__ const uint32_t c_inv_pio4[24] = {0xa2, 0xa2f9, 0xa2f983, 0xa2f9836e, 0xf9836e4e, 0x836e4e44, 0x6e4e4415, 0x4e441529, 0x441529fc, 0x1529fc27,
	0x29fc2757, 0xfc2757d1, 0x2757d1f5, 0x57d1f534, 0xd1f534dd, 0xf534ddc0, 0x34ddc0db, 0xddc0db62, 0xc0db6295, 0xdb629599, 0x6295993c, 0x95993c43, 0x993c4390,
	0x3c439041};

__device__ __forceinline__ double sc_reduce_large(uint32_t xi, int &n_out) {
	const uint32_t *arr = &c_inv_pio4[(xi >> 26) & 15];
	const int shift = (xi >> 23) & 7;
	xi = (xi & 0xffffffu) | 0x800000u;
	xi <<= shift;
	uint64_t res0 = uint32_t(xi * arr[0]);
	const uint64_t res1 = uint64_t(xi) * arr[4];
	const uint64_t res2 = uint64_t(xi) * arr[8];
	res0 = (res2 >> 32) | (res0 << 32);
	res0 += res1;
	const uint64_t n = (res0 + (1ULL << 61)) >> 62;
	res0 -= n << 62;
	n_out = int(n);
	return __dmul_rn(double(int64_t(res0)), 0x1.921FB54442D18p-62);
}

// The two polynomials of sincosf.h on the reduced argument, each rounded once to float. Every operation is sign-symmetric
// under round-to-nearest, so sin_poly(-x) == -sin_poly(x) and the negated-coefficient cosine table of glibc (__sincosf_table[1])
// gives exactly -cos_poly(x2): quadrant signs are applied to the rounded floats instead of evaluating four variants.
// Coefficients in constant memory: DFMA/DMUL take them as c[bank][offset] operands (an immediate double costs two UMOVs per use).
__constant__ double c_sc[11] = {
	-0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7, -0x1.555545995a603p-3,                             // S3 S2 S1
	0x1.99343027bf8c3p-16, -0x1.6c087e89a359dp-10, -0x1.ffffffd0c621cp-2, 0x1p0, 0x1.55553e1068f19p-5, // C4 C3 C1 C0 C2
	0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, 0x1.921FB54442D18p-62};                             // 2^24*2/pi, pi/2, pi/2 * 2^-62
__device__ __forceinline__ float sc_sin_poly(double x, double x2) {
	const double x3 = __dmul_rn(x, x2);
	const double s1 = fma(x2, c_sc[0], c_sc[1]);
	const double x7 = __dmul_rn(x3, x2);
	const double s = fma(x3, c_sc[2], x);
	return __double2float_rn(fma(x7, s1, s));
}
__device__ __forceinline__ float sc_cos_poly(double x2) {
	const double x4 = __dmul_rn(x2, x2);
	const double c2 = fma(x2, c_sc[3], c_sc[4]);
	const double c1 = fma(x2, c_sc[5], c_sc[6]);
	const double x6 = __dmul_rn(x4, x2);
	const double c = fma(x4, c_sc[7], c1);
	return __double2float_rn(fma(x6, c2, c));
}

// |y| >= 120, inf, nan: rare (Euler angles are within a few turns), kept out of line
__device__ __noinline__ float2 sincos_libm_slow(float y) {
	if (abstop12(y) >= 0x7f8u)
		return make_float2(y - y, y - y); // inf/nan -> nan
	const uint32_t xi = __float_as_uint(y);
	int n;
	const double x = sc_reduce_large(xi, n);
	const int m = n + int(xi >> 31); // s_sinf.c / s_cosf.c large path: sign and table are keyed on n + sign, the polynomial on n
	const double x2 = __dmul_rn(x, x);
	const float sp = sc_sin_poly(x, x2), cp = sc_cos_poly(x2);
	const float sgn_sp = ((m & 3) == 1 || (m & 3) == 2) ? -sp : sp; // sign[] = {1,-1,-1,1}
	const float neg_cp = (m & 2) ? -cp : cp;                        // __sincosf_table[1] = negated cosine coefficients
	return make_float2((n & 1) ? neg_cp : sgn_sp, (n & 1) ? sgn_sp : neg_cp);
}

// sinf(y) and cosf(y) exactly as glibc computes them (each is an independent libm call in the reference). One branch-free path
// for every |y| < 120: the |y| < pi/4 case of s_sinf.c is the general reduction with n = 0 (x is left untouched), and the
// |y| < 2^-12 shortcut is a select. Proven equal to libm for all 2^32 inputs on the host: oracle/exhaustive_sincos_merged.c.
__device__ __forceinline__ void sincos_libm(float y, float &s_out, float &c_out) {
	const uint32_t top = abstop12(y);
	if (top >= 0x42fu) {
		const float2 sc = sincos_libm_slow(y);
		s_out = sc.x, c_out = sc.y;
		return;
	}
	const double x = double(y);
	const double r = __dmul_rn(x, c_sc[8]);
	const int n = (__double2int_rz(r) + 0x800000) >> 24;
	const double xr = fma(-double(n), c_sc[9], x);
	const double x2 = __dmul_rn(xr, xr);
	const float sp = sc_sin_poly(xr, x2), cp = sc_cos_poly(x2);
	float s = (n & 1) ? cp : sp, c = (n & 1) ? sp : cp;
	s = (n & 2) ? -s : s;
	c = ((n + 1) & 2) ? -c : c;
	const bool tiny = top < 0x398u; // |y| < 2^-12
	s_out = tiny ? y : s;
	c_out = tiny ? 1.0f : c;
}

// The three Euler angles of a node at once: ONE range test and one branch instead of three (|y| >= 120, inf and nan of any angle send
// all three through the general path below), and the three fast evaluations form one straight-line block, so the polynomial
// coefficients are fetched once. Results are those of three sincos_libm calls.
__device__ __noinline__ void sincos3_libm_slow(float y0, float y1, float y2, float *sc6) {
	sincos_libm(y0, sc6[0], sc6[1]);
	sincos_libm(y1, sc6[2], sc6[3]);
	sincos_libm(y2, sc6[4], sc6[5]);
}
__device__ __forceinline__ void sincos_fast(float y, float &s_out, float &c_out) { // |y| < 120 (sincos_libm without the range test)
	const double x = double(y);
	const double r = __dmul_rn(x, c_sc[8]);
	const int n = (__double2int_rz(r) + 0x800000) >> 24;
	const double xr = fma(-double(n), c_sc[9], x);
	const double x2 = __dmul_rn(xr, xr);
	const float sp = sc_sin_poly(xr, x2), cp = sc_cos_poly(x2);
	float s = (n & 1) ? cp : sp, c = (n & 1) ? sp : cp;
	s = (n & 2) ? -s : s;
	c = ((n + 1) & 2) ? -c : c;
	const bool tiny = abstop12(y) < 0x398u; // |y| < 2^-12
	s_out = tiny ? y : s;
	c_out = tiny ? 1.0f : c;
}
__device__ __forceinline__ void sincos3_libm(float y0, float y1, float y2, float &s0, float &c0, float &s1, float &c1, float &s2, float &c2) {
	const uint32_t big = max(max(__float_as_uint(y0) & 0x7fffffffu, __float_as_uint(y1) & 0x7fffffffu), __float_as_uint(y2) & 0x7fffffffu);
	if (big >= 0x42f00000u) { // 120.0f
		float sc[6];
		sincos3_libm_slow(y0, y1, y2, sc);
		s0 = sc[0], c0 = sc[1], s1 = sc[2], c1 = sc[3], s2 = sc[4], c2 = sc[5];
		return;
	}
	sincos_fast(y0, s0, c0);
	sincos_fast(y1, s1, c1);
	sincos_fast(y2, s2, c2);
}

// ---- TransformationMat4(pos, euler, scale) ------------------------------------------------------------------------
// foundation/matrix4.cpp:176-183 -> RotationMat3(e, RO_YXZ) foundation/matrix3.cpp:180-260. For euler_order_yxz the unpacked
// axes are i=1 j=0 k=2 (odd, no repeat); the left-handed and the odd negations cancel exactly, leaving
// angle_0 = e.y, angle_1 = e.x, angle_2 = e.z.
__device__ __forceinline__ Mat4 transformation_mat4(float px, float py, float pz, float ex, float ey, float ez, float sx, float sy, float sz) {
	float s0, c0, s1, c1, s2, c2;
	sincos_libm(ey, s0, c0);
	sincos_libm(ex, s1, c1);
	sincos_libm(ez, s2, c2);
	// products in pairs that share a factor (same factors, same roundings as the scalar form)
	const f32x2_t s2c2 = pk2(s2, c2);
	float s0s2, s0c2, c0s2, c0c2, r10, r11, r02, r22;
	upk2(mul2(pk2(s0, s0), s2c2), s0s2, s0c2);
	upk2(mul2(pk2(c0, c0), s2c2), c0s2, c0c2);
	upk2(mul2(pk2(c1, c1), s2c2), r10, r11);
	upk2(mul2(pk2(c1, c1), pk2(s0, c0)), r02, r22);
	float a01, a00, a21, a20;
	upk2(mul2(pk2(s1, s1), pk2(s0c2, s0s2)), a01, a00);
	upk2(mul2(pk2(s1, s1), pk2(c0c2, c0s2)), a21, a20);
	const float r12 = -s1;
	const float r01 = a01 - c0s2, r00 = a00 + c0c2;
	const float r21 = a21 + s0s2, r20 = a20 - s0c2;
	Mat4 o;
	upk2(mul2(pk2(sx, sx), pk2(r00, r10)), o.m[0][0], o.m[1][0]);
	upk2(mul2(pk2(sy, sy), pk2(r01, r11)), o.m[0][1], o.m[1][1]);
	upk2(mul2(pk2(sz, sz), pk2(r02, r12)), o.m[0][2], o.m[1][2]);
	upk2(mul2(pk2(r20, r21), pk2(sx, sy)), o.m[2][0], o.m[2][1]);
	o.m[2][2] = r22 * sz;
	o.m[0][3] = px, o.m[1][3] = py, o.m[2][3] = pz;
	return o;
}

// operator*(Mat4, Mat4) foundation/matrix4.cpp:31-44 — left-to-right sums, translation column adds a[r][3] last
__device__ __forceinline__ Mat4 mat4_mul(const Mat4 &a, const Mat4 &b) {
	Mat4 o;
	const f32x2_t b00 = pk2(b.m[0][0], b.m[0][1]), b02 = pk2(b.m[0][2], b.m[0][3]);
	const f32x2_t b10 = pk2(b.m[1][0], b.m[1][1]), b12 = pk2(b.m[1][2], b.m[1][3]);
	const f32x2_t b20 = pk2(b.m[2][0], b.m[2][1]), b22 = pk2(b.m[2][2], b.m[2][3]);
#pragma unroll
	for (int r = 0; r < 3; ++r) {
		const f32x2_t a0 = pk2(a.m[r][0], a.m[r][0]), a1 = pk2(a.m[r][1], a.m[r][1]), a2 = pk2(a.m[r][2], a.m[r][2]);
		float p0, p1, p2, p3, q0, q1, q2, q3, t0, t1, t2, t3;
		upk2(mul2(a0, b00), p0, p1), upk2(mul2(a0, b02), p2, p3);
		upk2(mul2(a1, b10), q0, q1), upk2(mul2(a1, b12), q2, q3);
		upk2(mul2(a2, b20), t0, t1), upk2(mul2(a2, b22), t2, t3);
		o.m[r][0] = p0 + q0 + t0;
		o.m[r][1] = p1 + q1 + t1;
		o.m[r][2] = p2 + q2 + t2;
		o.m[r][3] = p3 + q3 + t3 + a.m[r][3];
	}
	return o;
}

// ---- row-pair form (level kernels) ------------------------------------------------------------------------------------------------
// The same two functions with the LOCAL matrix kept as the register pairs mul.f32x2 consumes — a[i] = (m[i][0], m[i][1]),
// b[i] = (m[i][2], m[i][3]) — so that no pair has to be assembled from scalars of two other pairs (profiles r02a: a third of the
// packed products' gain was going into MOVs). Same factors, same roundings, same sums as the scalar forms above.
struct Mat4p {
	f32x2_t a[3], b[3];
};
__device__ __forceinline__ Mat4p transformation_mat4p(float px, float py, float pz, float ex, float ey, float ez, float sx, float sy, float sz) {
	float s0, c0, s1, c1, s2, c2;
	sincos3_libm(ey, ex, ez, s0, c0, s1, c1, s2, c2);
	const f32x2_t s2c2 = pk2(s2, c2), sxy = pk2(sx, sy);
	const f32x2_t s0x = mul2(pk2(s0, s0), s2c2); // (s0s2, s0c2)
	const f32x2_t c0x = mul2(pk2(c0, c0), s2c2); // (c0s2, c0c2)
	float s0s2, s0c2, c0s2, c0c2, a00, a01, a20, a21, r02, r22;
	upk2(s0x, s0s2, s0c2);
	upk2(c0x, c0s2, c0c2);
	upk2(mul2(pk2(s1, s1), s0x), a00, a01);
	upk2(mul2(pk2(s1, s1), c0x), a20, a21);
	upk2(mul2(pk2(c1, c1), pk2(s0, c0)), r02, r22);
	Mat4p o;
	o.a[0] = mul2(pk2(a00 + c0c2, a01 - c0s2), sxy);
	o.a[1] = mul2(mul2(pk2(c1, c1), s2c2), sxy);
	o.a[2] = mul2(pk2(a20 - s0c2, a21 + s0s2), sxy);
	o.b[0] = pk2(r02 * sz, px);
	o.b[1] = pk2(-s1 * sz, py);
	o.b[2] = pk2(r22 * sz, pz);
	return o;
}
__device__ __forceinline__ Mat4 mat4_mul_p(const Mat4 &a, const Mat4p &b) {
	Mat4 o;
#pragma unroll
	for (int r = 0; r < 3; ++r) {
		const f32x2_t a0 = pk2(a.m[r][0], a.m[r][0]), a1 = pk2(a.m[r][1], a.m[r][1]), a2 = pk2(a.m[r][2], a.m[r][2]);
		float p0, p1, p2, p3, q0, q1, q2, q3, t0, t1, t2, t3;
		upk2(mul2(a0, b.a[0]), p0, p1), upk2(mul2(a0, b.b[0]), p2, p3);
		upk2(mul2(a1, b.a[1]), q0, q1), upk2(mul2(a1, b.b[1]), q2, q3);
		upk2(mul2(a2, b.a[2]), t0, t1), upk2(mul2(a2, b.b[2]), t2, t3);
		o.m[r][0] = p0 + q0 + t0;
		o.m[r][1] = p1 + q1 + t1;
		o.m[r][2] = p2 + q2 + t2;
		o.m[r][3] = p3 + q3 + t3 + a.m[r][3];
	}
	return o;
}
__device__ __forceinline__ Mat4 unpack_mat4(const Mat4p &p) {
	Mat4 o;
#pragma unroll
	for (int r = 0; r < 3; ++r)
		upk2(p.a[r], o.m[r][0], o.m[r][1]), upk2(p.b[r], o.m[r][2], o.m[r][3]);
	return o;
}

__device__ __forceinline__ Mat4 load_mat4(const float *__restrict__ p) { // 48 B, 16-B aligned
	const float4 *q = reinterpret_cast<const float4 *>(p);
	const float4 a = q[0], b = q[1], c = q[2];
	Mat4 o;
	o.m[0][0] = a.x, o.m[0][1] = a.y, o.m[0][2] = a.z, o.m[0][3] = a.w;
	o.m[1][0] = b.x, o.m[1][1] = b.y, o.m[1][2] = b.z, o.m[1][3] = b.w;
	o.m[2][0] = c.x, o.m[2][1] = c.y, o.m[2][2] = c.z, o.m[2][3] = c.w;
	return o;
}
__device__ __forceinline__ void store_mat4(float *__restrict__ p, const Mat4 &o) {
	float4 *q = reinterpret_cast<float4 *>(p);
	q[0] = make_float4(o.m[0][0], o.m[0][1], o.m[0][2], o.m[0][3]);
	q[1] = make_float4(o.m[1][0], o.m[1][1], o.m[1][2], o.m[1][3]);
	q[2] = make_float4(o.m[2][0], o.m[2][1], o.m[2][2], o.m[2][3]);
}

// ---- culling predicate -----------------------------------------------------------------------------------------------
// operator*(Mat4, MinMax) foundation/minmax.cpp:226-244 (Arvo) followed by TestVisibility(Frustum, MinMax)
// foundation/frustum.cpp:59-79. Returns true when the entry SURVIVES CullModelDisplayLists (engine/render_pipeline.cpp:2098-2104),
// i.e. visibility != V_Outside.
struct Aabb {
	float mn[3], mx[3];
};

// The device MinMax table is stored interleaved: {mn.x, mx.x, mn.y, mx.y, mn.z, mx.z} (hgb_set_bounds converts), so (mn[i], mx[i]) is
// one 8-byte pair: both products of a matrix element are one mul2, and (mn, mx) += (lo, hi) is one add2 (its operands come from
// selects, never straight from a product).
__device__ __forceinline__ Aabb minmax_transform(const Mat4 &m, const float *__restrict__ b6) {
	const float2 *bp = reinterpret_cast<const float2 *>(b6);
	const float2 bx = bp[0], by = bp[1], bz = bp[2];
	const f32x2_t bb[3] = {pk2(bx.x, bx.y), pk2(by.x, by.y), pk2(bz.x, bz.y)};
	Aabb o;
#pragma unroll
	for (int j = 0; j < 3; ++j) {
		f32x2_t acc = pk2(m.m[j][3], m.m[j][3]);
#pragma unroll
		for (int i = 0; i < 3; ++i) {
			float a, b;
			upk2(mul2(pk2(m.m[j][i], m.m[j][i]), bb[i]), a, b);
			const bool lt = a < b;
			acc = add2(acc, pk2(lt ? a : b, lt ? b : a));
		}
		upk2(acc, o.mn[j], o.mx[j]);
	}
	return o;
}

// planes are stored pre-conditioned: {x, y, z, -2*w} (the reference computes plane_d_x2 = -plane.w * 2.f per test; exact either way)
__device__ __forceinline__ bool survives_frustum(const float4 *__restrict__ planes, const Aabb &bb) {
	const float cx = bb.mn[0] + bb.mx[0], cy = bb.mn[1] + bb.mx[1], cz = bb.mn[2] + bb.mx[2];
	const float ex = bb.mx[0] - bb.mn[0], ey = bb.mx[1] - bb.mn[1], ez = bb.mx[2] - bb.mn[2];
	bool inside = true;
#pragma unroll
	for (int n = 0; n < 6; ++n) {
		const float4 p = planes[n];
		const float d = p.x * cx + p.y * cy + p.z * cz;
		const float r = fabsf(p.x) * ex + fabsf(p.y) * ey + fabsf(p.z) * ez;
		inside = inside && !(d - r > p.w);
	}
	return inside;
}

// Same test with the planes expanded in shared memory to two float4 each: {x, |x|, y, |y|} {z, |z|, -2w, 0}, so that the products of
// d and r pair up: (x*cx, |x|*ex) is one mul2. Sums stay scalar, left to right as in the reference.
__device__ __forceinline__ void expand_plane(const float4 p, float4 &a, float4 &b) {
	a = make_float4(p.x, fabsf(p.x), p.y, fabsf(p.y));
	b = make_float4(p.z, fabsf(p.z), p.w, 0.f);
}
__device__ __forceinline__ bool survives_frustum_x(const float4 *__restrict__ planes2, const Aabb &bb) {
	const f32x2_t cex = pk2(bb.mn[0] + bb.mx[0], bb.mx[0] - bb.mn[0]);
	const f32x2_t cey = pk2(bb.mn[1] + bb.mx[1], bb.mx[1] - bb.mn[1]);
	const f32x2_t cez = pk2(bb.mn[2] + bb.mx[2], bb.mx[2] - bb.mn[2]);
	uint32_t outside = 0; // all-ones per failed plane, OR-ed: no early out, no select chain
#pragma unroll
	for (int n = 0; n < 6; ++n) {
		const float4 a = planes2[2 * n], b = planes2[2 * n + 1];
		float dx, rx, dy, ry, dz, rz;
		upk2(mul2(pk2(a.x, a.y), cex), dx, rx);
		upk2(mul2(pk2(a.z, a.w), cey), dy, ry);
		upk2(mul2(pk2(b.x, b.y), cez), dz, rz);
		const float d = dx + dy + dz;
		const float r = rx + ry + rz;
		uint32_t o;
		asm("set.gt.u32.f32 %0, %1, %2;" : "=r"(o) : "f"(d - r), "f"(b.z));
		outside |= o;
	}
	return outside == 0;
}

} // namespace hgb
