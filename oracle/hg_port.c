/* TEST INFRASTRUCTURE ONLY — see hg_port.h. Plain-C restatement of the reference's scene-update hot path.
 * Build: gcc -O2 -mfma -ffp-contract=off  (single-precision ops stay unfused, in source order; the only fused
 * operations are the explicit fma() calls of the libm sinf/cosf restatement). */
#include "hg_port.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- sinf / cosf ---------------------------------------------------------------------------------------------
 * The reference calls libm sinf/cosf (matrix3.cpp:199-204). glibc 2.39 implements them in
 * sysdeps/ieee754/flt-32/{s_sinf.c,s_cosf.c,sincosf.h,sincosf_data.c} (ARM optimized-routines algorithm): reduce
 * in double, evaluate a double polynomial, round once to float. On FMA-capable x86-64 the ifunc picks the -mfma
 * build, where every a*b+c below is one fused op. Restated from the published algorithm; constants are the
 * published table. Verified against libm over all 2^32 inputs (exhaustive_sincos.c). */
typedef struct {
	double sign[4];
	double hpi_inv, hpi, c0, c1, c2, c3, c4, s1, s2, s3;
} sincos_tab;

static const sincos_tab SC[2] = {
	{{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, 0x1p0, -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10,
		0x1.99343027bf8c3p-16, -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13},
	{{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, -0x1p0, 0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10,
		-0x1.99343027bf8c3p-16, -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13}};

/* 4/pi as a 24 x 32-bit sliding window, for |x| >= 120 */
static const uint32_t INV_PIO4[24] = {0xa2, 0xa2f9, 0xa2f983, 0xa2f9836e, 0xf9836e4e, 0x836e4e44, 0x6e4e4415, 0x4e441529, 0x441529fc, 0x1529fc27, 0x29fc2757,
	0xfc2757d1, 0x2757d1f5, 0x57d1f534, 0xd1f534dd, 0xf534ddc0, 0x34ddc0db, 0xddc0db62, 0xc0db6295, 0xdb629599, 0x6295993c, 0x95993c43, 0x993c4390, 0x3c439041};

static inline uint32_t f2u(float f) {
	uint32_t u;
	memcpy(&u, &f, 4);
	return u;
}
static inline uint32_t abstop12(float x) { return (f2u(x) >> 20) & 0x7ff; }

static inline double reduce_fast(double x, const sincos_tab *p, int *np) {
	const double r = x * p->hpi_inv;
	const int n = ((int32_t)r + 0x800000) >> 24;
	*np = n;
	return fma(-(double)n, p->hpi, x);
}

static inline double reduce_large(uint32_t xi, int *np) {
	const uint32_t *arr = &INV_PIO4[(xi >> 26) & 15];
	const int shift = (xi >> 23) & 7;
	uint64_t n, res0, res1, res2;
	xi = (xi & 0xffffff) | 0x800000;
	xi <<= shift;
	res0 = xi * arr[0];
	res1 = (uint64_t)xi * arr[4];
	res2 = (uint64_t)xi * arr[8];
	res0 = (res2 >> 32) | (res0 << 32);
	res0 += res1;
	n = (res0 + (1ULL << 61)) >> 62;
	res0 -= n << 62;
	*np = (int)n;
	return (double)(int64_t)res0 * 0x1.921FB54442D18p-62;
}

static inline float sincos_poly(double x, double x2, const sincos_tab *p, int n) {
	if ((n & 1) == 0) {
		const double x3 = x * x2;
		const double s1 = fma(x2, p->s3, p->s2);
		const double x7 = x3 * x2;
		const double s = fma(x3, p->s1, x);
		return (float)fma(x7, s1, s);
	} else {
		const double x4 = x2 * x2;
		const double c2 = fma(x2, p->c4, p->c3);
		const double c1 = fma(x2, p->c1, p->c0);
		const double x6 = x4 * x2;
		const double c = fma(x4, p->c2, c1);
		return (float)fma(x6, c2, c);
	}
}

static float sincos_eval(float y, int want_cos) {
	double x = y, s;
	int n;
	const sincos_tab *p = &SC[0];
	if (abstop12(y) < abstop12(0x1.921FB6p-1f)) { /* |y| < pi/4 */
		if (abstop12(y) < abstop12(0x1p-12f))
			return want_cos ? 1.0f : y;
		return sincos_poly(x, x * x, p, want_cos);
	} else if (abstop12(y) < abstop12(120.0f)) {
		x = reduce_fast(x, p, &n);
		s = p->sign[(n + want_cos) & 3];
		if (n & 2)
			p = &SC[1];
		return sincos_poly(x * s, x * x, p, n ^ want_cos);
	} else if (abstop12(y) < abstop12(INFINITY)) {
		const uint32_t xi = f2u(y);
		const int sign = xi >> 31;
		x = reduce_large(xi, &n);
		s = p->sign[(n + sign) & 3];
		if ((n + sign) & 2)
			p = &SC[1];
		return sincos_poly(x * s, x * x, p, n ^ want_cos);
	}
	return y - y; /* inf/nan -> nan */
}

float hgp_sinf(float x) { return sincos_eval(x, 0); }
float hgp_cosf(float x) { return sincos_eval(x, 1); }

/* ---- matrices -------------------------------------------------------------------------------------------------- */
#define M(m, r, c) ((m)[(r) * 4 + (c)])

/* matrix3.cpp:180-234 specialised for euler_order_yxz (0x06): i=1 j=0 k=2, odd, no repeat. The left-handed negation
 * (:183-185) and the odd negation (:193-197) cancel exactly, so angle_0 = e.y, angle_1 = e.x, angle_2 = e.z.
 * R[r][c] below is hg::Mat3::m[r][c]. */
static void rotation_yxz(float ex, float ey, float ez, float R[3][3]) {
	const float s0 = hgp_sinf(ey), c0 = hgp_cosf(ey);
	const float s1 = hgp_sinf(ex), c1 = hgp_cosf(ex);
	const float s2 = hgp_sinf(ez), c2 = hgp_cosf(ez);
	const float s0s2 = s0 * s2, s0c2 = s0 * c2, c0s2 = c0 * s2, c0c2 = c0 * c2;
	R[1][1] = c1 * c2;
	R[1][0] = c1 * s2;
	R[1][2] = -s1;
	R[0][1] = s1 * s0c2 - c0s2;
	R[0][0] = s1 * s0s2 + c0c2;
	R[0][2] = s0 * c1;
	R[2][1] = s1 * c0c2 + s0s2;
	R[2][0] = s1 * c0s2 - s0c2;
	R[2][2] = c0 * c1;
}

void hgp_transformation_mat4(const float *t, float *o) {
	float R[3][3];
	rotation_yxz(t[3], t[4], t[5], R);
	for (int r = 0; r < 3; ++r) {
		M(o, r, 0) = R[r][0] * t[6];
		M(o, r, 1) = R[r][1] * t[7];
		M(o, r, 2) = R[r][2] * t[8];
		M(o, r, 3) = t[r];
	}
}

void hgp_mat4_mul(const float *a, const float *b, float *out) {
	float o[12];
	for (int r = 0; r < 3; ++r) {
		for (int c = 0; c < 3; ++c)
			M(o, r, c) = M(a, r, 0) * M(b, 0, c) + M(a, r, 1) * M(b, 1, c) + M(a, r, 2) * M(b, 2, c);
		M(o, r, 3) = M(a, r, 0) * M(b, 0, 3) + M(a, r, 1) * M(b, 1, 3) + M(a, r, 2) * M(b, 2, 3) + M(a, r, 3);
	}
	memcpy(out, o, sizeof(o));
}

void hgp_inverse_fast(const float *m, float *out) {
	float o[12], S[3];
	for (int c = 0; c < 3; ++c)
		S[c] = 1.F / (M(m, 0, c) * M(m, 0, c) + M(m, 1, c) * M(m, 1, c) + M(m, 2, c) * M(m, 2, c));
	for (int r = 0; r < 3; ++r)
		for (int c = 0; c < 3; ++c)
			M(o, r, c) = M(m, c, r) * S[r];
	const float Tx = -M(m, 0, 3), Ty = -M(m, 1, 3), Tz = -M(m, 2, 3);
	for (int r = 0; r < 3; ++r)
		M(o, r, 3) = M(o, r, 0) * Tx + M(o, r, 1) * Ty + M(o, r, 2) * Tz;
	memcpy(out, o, sizeof(o));
}

void hgp_minmax_transform(const float *m, const float *mm, float *out) {
	float mn[3], mx[3];
	for (int j = 0; j < 3; ++j) {
		mn[j] = mx[j] = M(m, j, 3);
		for (int i = 0; i < 3; ++i) {
			const float a = M(m, j, i) * mm[i], b = M(m, j, i) * mm[3 + i];
			if (a < b) {
				mn[j] += a;
				mx[j] += b;
			} else {
				mn[j] += b;
				mx[j] += a;
			}
		}
	}
	memcpy(out, mn, 12);
	memcpy(out + 3, mx, 12);
}

int hgp_test_visibility(const float *f, const float *mm) {
	const float cx = mm[0] + mm[3], cy = mm[1] + mm[4], cz = mm[2] + mm[5];
	const float ex = mm[3] - mm[0], ey = mm[4] - mm[1], ez = mm[5] - mm[2];
	int vis = 1;
	for (int n = 0; n < 6; ++n) {
		const float *p = f + n * 4;
		const float d = p[0] * cx + p[1] * cy + p[2] * cz;
		const float r = fabsf(p[0]) * ex + fabsf(p[1]) * ey + fabsf(p[2]) * ez;
		const float plane_d_x2 = -p[3] * 2.f;
		if (d - r > plane_d_x2)
			return 0;
		if (d + r > plane_d_x2)
			vis = 2;
	}
	return vis;
}

void hgp_to_bgfx(const float *m, float *o) {
	for (int c = 0; c < 4; ++c) {
		o[c * 4 + 0] = M(m, 0, c);
		o[c * 4 + 1] = M(m, 1, c);
		o[c * 4 + 2] = M(m, 2, c);
		o[c * 4 + 3] = c == 3 ? 1.f : 0.f;
	}
}

/* ---- view frustum (host side, once per view) ------------------------------------------------------------------- */
static void normalize_reverse_plane(const float *p, float *o) { /* frustum.cpp:12-15 + Reverse (vector4.h:160) */
	const float l = sqrtf(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
	for (int i = 0; i < 4; ++i)
		o[i] = -(p[i] / l);
}

void hgp_make_perspective_frustum(float znear, float zfar, float fov, float ar_x, float ar_y, const float *world, float *out24) {
	/* FovToZoomFactor projection.cpp:19 — Clamp(fov, Deg(0.1f), Deg(179.9f)); Tan() is double tan (math.cpp:93) */
	const float lo = 0.1f / 180.f * 3.1415926535f, hi = 179.9f / 180.f * 3.1415926535f;
	const float cf = fov < lo ? lo : (fov > hi ? hi : fov);
	const float zoom = 1.f / (float)tan((double)(cf * 0.5f));
	/* ComputePerspectiveProjectionMatrix projection.cpp:29-37, NDC {origin_bottom_left=false, homogeneous_depth=false} */
	const float qA = zfar / (zfar - znear), qB = -qA * znear;
	const float X[4] = {zoom / ar_x, 0, 0, 0}, Y[4] = {0, zoom / ar_y, 0, 0}, Z[4] = {0, 0, qA, qB}, W[4] = {0, 0, 1, 0};
	float local[6][4], tmp[4];
	const float *rows[3] = {Y, X, Z};
	for (int k = 0; k < 3; ++k) { /* W+Y, W-Y, W+X, W-X, W+Z, W-Z  (frustum.cpp:19-20) */
		for (int i = 0; i < 4; ++i)
			tmp[i] = W[i] + rows[k][i];
		normalize_reverse_plane(tmp, local[k * 2]);
		for (int i = 0; i < 4; ++i)
			tmp[i] = W[i] - rows[k][i];
		normalize_reverse_plane(tmp, local[k * 2 + 1]);
	}
	/* TransformFrustum frustum.cpp:25-28: iMt = Transpose(Mat44(InverseFast(m))); plane' = iMt * plane (matrix44.cpp:76-79) */
	float inv[12], I44[4][4];
	hgp_inverse_fast(world, inv);
	for (int r = 0; r < 3; ++r)
		for (int c = 0; c < 4; ++c)
			I44[r][c] = M(inv, r, c);
	I44[3][0] = I44[3][1] = I44[3][2] = 0.f;
	I44[3][3] = 1.f;
	for (int n = 0; n < 6; ++n) {
		const float *v = local[n];
		for (int r = 0; r < 4; ++r) /* iMt[r][c] = I44[c][r] */
			out24[n * 4 + r] = v[0] * I44[0][r] + v[1] * I44[1][r] + v[2] * I44[2][r] + v[3] * I44[3][r];
	}
}

/* ---- quaternion rotation tracks: ToEuler(Normalize(q)) (scene.cpp:2172) ---------------------------------------- */
void hgp_quat_to_euler(const float *q, float *out3) {
	float x = q[0], y = q[1], z = q[2], w = q[3];
	const float l = sqrtf(x * x + y * y + z * z + w * w); /* quaternion.cpp:54,57-60 */
	if (l) {
		x = x / l, y = y / l, z = z / l, w = w / l;
	} else {
		x = y = z = 0.f, w = 1.f;
	}
	/* ToMatrix3 quaternion.cpp:141-166; m[r][c] == hg::Mat3::m[r][c] */
	float m[3][3];
	const float sqw = w * w, sqx = x * x, sqy = y * y, sqz = z * z;
	const float invs = 1.f / (sqx + sqy + sqz + sqw);
	m[0][0] = (sqx - sqy - sqz + sqw) * invs;
	m[1][1] = (-sqx + sqy - sqz + sqw) * invs;
	m[2][2] = (-sqx - sqy + sqz + sqw) * invs;
	float tmp1 = x * y, tmp2 = z * w;
	m[1][0] = 2.f * (tmp1 + tmp2) * invs;
	m[0][1] = 2.f * (tmp1 - tmp2) * invs;
	tmp1 = x * z, tmp2 = y * w;
	m[2][0] = 2.f * (tmp1 - tmp2) * invs;
	m[0][2] = 2.f * (tmp1 + tmp2) * invs;
	tmp1 = y * z, tmp2 = x * w;
	m[2][1] = 2.f * (tmp1 + tmp2) * invs;
	m[1][2] = 2.f * (tmp1 - tmp2) * invs;
	/* ToEuler matrix3.cpp:91-137 for euler_order_yxz: i=1 j=0 k=2, odd, no repeat; the odd negation and the
	 * left-handed negation cancel. atan2 is the DOUBLE libm function (float args promoted), result narrowed to float. */
	const float tolerance = 1.1920929e-07f; /* numeric_limits<float>::epsilon() */
	const float s1 = -m[1][2];
	const float c1 = sqrtf(m[1][1] * m[1][1] + m[1][0] * m[1][0]);
	float a0, a1, a2;
	a1 = (float)atan2((double)s1, (double)c1);
	if (c1 > tolerance) {
		a0 = (float)atan2((double)m[0][2], (double)m[2][2]);
		a2 = (float)atan2((double)m[1][0], (double)m[1][1]);
	} else {
		a0 = 0.F;
		const float sign_s1 = s1 < 0.F ? -1.F : (s1 > 0.F ? 1.F : 0.F);
		a2 = (float)((double)(-sign_s1) * atan2((double)-m[2][0], (double)m[0][0]));
	}
	a0 = -(-a0), a1 = -(-a1), a2 = -(-a2);
	out3[0] = a1; /* ToEuler(m, yxz, euler.y, euler.x, euler.z) matrix3.cpp:152 */
	out3[1] = a0;
	out3[2] = a2;
}

/* ---- animation sampling ---------------------------------------------------------------------------------------- */
int hgp_get_interval_keys(int n, const int64_t *t, int64_t at, int *kf0, int *kf1) {
	int i = 0;
	for (; i < n; ++i)
		if (t[i] > at)
			break;
	if (i == 0) {
		*kf0 = 0;
		return 0;
	} else if (i == n) {
		*kf0 = i - 1;
		return 0;
	}
	*kf0 = i - 1;
	*kf1 = i;
	return 1;
}

static inline float time_to_sec_f(int64_t t) { return (float)((double)t / 1000000000.0); } /* time.h:13 */

int hgp_eval_vec3_track(int n, const int64_t *t, const float *v, const float *tension, const float *bias, int64_t at, float *out3) {
	if (n == 0)
		return 0;
	int kf1, kf2;
	if (hgp_get_interval_keys(n, t, at, &kf1, &kf2)) {
		const float u = time_to_sec_f(at - t[kf1]) / time_to_sec_f(t[kf2] - t[kf1]);
		const int kf0 = kf1 - 1 > 0 ? kf1 - 1 : 0, kf3 = kf2 + 1 < n - 1 ? kf2 + 1 : n - 1;
		const float ten = tension ? tension[kf1] : 0.f, bi = bias ? bias[kf1] : 0.f;
		/* HermiteInterpolate math.h:137-149, per component of Vec3 */
		const float t2 = u * u, t3 = t2 * u;
		const float a0 = 2.F * t3 - 3.F * t2 + 1.F;
		const float a1 = t3 - 2.F * t2 + u;
		const float a2 = t3 - t2;
		const float a3 = -2.F * t3 + 3.F * t2;
		for (int c = 0; c < 3; ++c) {
			const float y0 = v[kf0 * 3 + c], y1 = v[kf1 * 3 + c], y2 = v[kf2 * 3 + c], y3 = v[kf3 * 3 + c];
			float m0 = (y1 - y0) * (1.F + bi) * (1.F - ten) / 2.F;
			m0 += (y2 - y1) * (1.F - bi) * (1.F - ten) / 2.F;
			float m1 = (y2 - y1) * (1.F + bi) * (1.F - ten) / 2.F;
			m1 += (y3 - y2) * (1.F - bi) * (1.F - ten) / 2.F;
			out3[c] = y1 * a0 + m0 * a1 + m1 * a2 + y2 * a3;
		}
	} else {
		memcpy(out3, v + kf1 * 3, 12);
	}
	return 1;
}

int hgp_eval_quat_track(int n, const int64_t *t, const float *v, int64_t at, float *out4) {
	if (n == 0)
		return 0;
	int kf0, kf1;
	if (hgp_get_interval_keys(n, t, at, &kf0, &kf1)) {
		const float k = time_to_sec_f(at - t[kf0]) / time_to_sec_f(t[kf1] - t[kf0]);
		for (int c = 0; c < 4; ++c)
			out4[c] = v[kf0 * 4 + c] + (v[kf1 * 4 + c] - v[kf0 * 4 + c]) * k;
	} else {
		memcpy(out4, v + kf0 * 4, 16);
	}
	return 1;
}

int hgp_step_play_clock(int64_t *t, int64_t dt, int8_t t_scale, int loop_mode, int64_t t_start, int64_t t_end, int paused) {
	int done = 0;
	if (!paused)
		*t += (dt * t_scale) >> 4;
	if (loop_mode == 1) {
	} else if (loop_mode == 2) {
		if (t_scale >= 0) {
			while (*t >= t_end)
				*t -= t_end - t_start;
		} else {
			while (*t <= t_start)
				*t += t_end - t_start;
		}
	} else {
		if (t_scale >= 0) {
			if (*t >= t_end) {
				*t = t_end;
				done = 1;
			}
		} else if (*t <= t_start) {
			*t = t_start;
			done = 1;
		}
	}
	return done;
}

/* ---- scene level ------------------------------------------------------------------------------------------------ */
void hgp_compute_world_matrices(int n, const float *trs, const int32_t *parent, const uint8_t *used, const uint8_t *preset, float *worlds) {
	uint8_t *updated = (uint8_t *)calloc(n > 0 ? n : 1, 1);
	int32_t *stack = (int32_t *)malloc(sizeof(int32_t) * (n > 0 ? n : 1));
	if (preset)
		for (int i = 0; i < n; ++i)
			updated[i] = preset[i] ? 1 : 0;
	for (int i = 0; i < n; ++i) {
		if (used && !used[i])
			continue;
		/* the reference recurses (scene.cpp:263-277); an explicit stack gives the same memoised order */
		int sp = 0;
		for (int k = i; k >= 0 && !updated[k]; k = parent[k])
			stack[sp++] = k;
		while (sp > 0) {
			const int k = stack[--sp];
			float local[12];
			hgp_transformation_mat4(trs + (size_t)k * 9, local);
			if (parent[k] >= 0)
				hgp_mat4_mul(worlds + (size_t)parent[k] * 12, local, worlds + (size_t)k * 12);
			else
				memcpy(worlds + (size_t)k * 12, local, 48);
			updated[k] = 1;
		}
	}
	free(updated);
	free(stack);
}

int hgp_cull(int n, const uint32_t *mtx_idx, const uint32_t *bounds_idx, const float *bounds6, const float *worlds, const float *frustum,
	uint32_t *out_visible) {
	int k = 0;
	for (int e = 0; e < n; ++e) {
		float mm[6];
		hgp_minmax_transform(worlds + (size_t)mtx_idx[e] * 12, bounds6 + (size_t)bounds_idx[e] * 6, mm);
		if (hgp_test_visibility(frustum, mm) != 0)
			out_visible[k++] = (uint32_t)e;
	}
	return k;
}

void hgp_bone_matrices(int n, const uint32_t *mtx_idx, const uint32_t *bind_idx, const float *worlds, const float *bind_pose, float *out16) {
	for (int i = 0; i < n; ++i) {
		float m[12];
		hgp_mat4_mul(worlds + (size_t)mtx_idx[i] * 12, bind_pose + (size_t)bind_idx[i] * 12, m);
		hgp_to_bgfx(m, out16 + (size_t)i * 16);
	}
}
