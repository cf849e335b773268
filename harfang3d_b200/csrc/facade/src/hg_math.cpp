// hg::Scene facade — edit-time math helpers (camera matrices, projections, frusta) with the reference's operation order, so that
// a frustum built here has the same bits as one built by the reference (checked against tests/golden/math.npz by tests/test_facade.py).
// Build with -ffp-contract=off. The per-frame arithmetic of the path is NOT here: it runs on the GPU (csrc/hgb_math.cuh).
#include "foundation/hg_math_types.h"
#include "foundation/log.h"
#include <cstdio>

namespace hg {

const Vec3 Vec3::Zero(0, 0, 0), Vec3::One(1, 1, 1);
const Mat4 Mat4::Identity(Mat4::raw_tag{}, 1.f), Mat4::Zero(Mat4::raw_tag{}, 0.f);

// ---- log (foundation/log.h) ----
static int g_log_mask = LL_All;
static void (*g_log_hook)(const char *, int, const char *, void *) = nullptr;
static void *g_log_user = nullptr;
void set_log_level(int mask) { g_log_mask = mask; }
void set_log_hook(void (*on_log)(const char *, int, const char *, void *), void *user) { g_log_hook = on_log, g_log_user = user; }
static void emit(const char *msg, int level, const char *tag) {
	if (!(g_log_mask & level))
		return;
	if (g_log_hook)
		g_log_hook(msg, level, nullptr, g_log_user);
	else
		fprintf(stderr, "%s%s\n", tag, msg);
}
void log(const char *msg) { emit(msg, LL_Normal, ""); }
void warn(const char *msg) { emit(msg, LL_Warning, "WARNING: "); }
void error(const char *msg) { emit(msg, LL_Error, "ERROR: "); }

// ---- Mat4 ----
Mat4 operator*(const Mat4 &a, const Mat4 &b) { // affine product, sums left to right, translation added last (matrix4.cpp:31-44)
	Mat4 o;
	for (int r = 0; r < 3; ++r) {
		for (int c = 0; c < 3; ++c)
			o.m[r][c] = a.m[r][0] * b.m[0][c] + a.m[r][1] * b.m[1][c] + a.m[r][2] * b.m[2][c];
		o.m[r][3] = a.m[r][0] * b.m[0][3] + a.m[r][1] * b.m[1][3] + a.m[r][2] * b.m[2][3] + a.m[r][3];
	}
	return o;
}

Mat4 InverseFast(const Mat4 &m) { // inverse of a scaled rotation + translation (matrix4.cpp:293-303): transpose / squared column length
	float s[3];
	for (int c = 0; c < 3; ++c)
		s[c] = 1.f / (m.m[0][c] * m.m[0][c] + m.m[1][c] * m.m[1][c] + m.m[2][c] * m.m[2][c]);
	Mat4 o;
	for (int r = 0; r < 3; ++r)
		for (int c = 0; c < 3; ++c)
			o.m[r][c] = m.m[c][r] * s[r];
	const float tx = -m.m[0][3], ty = -m.m[1][3], tz = -m.m[2][3];
	for (int r = 0; r < 3; ++r)
		o.m[r][3] = o.m[r][0] * tx + o.m[r][1] * ty + o.m[r][2] * tz;
	return o;
}

Mat4 TranslationMat4(const Vec3 &p) {
	Mat4 o;
	o.m[0][3] = p.x, o.m[1][3] = p.y, o.m[2][3] = p.z;
	return o;
}

// Euler YXZ (the reference's default order, rotation_order.h:17): angle_0 = e.y, angle_1 = e.x, angle_2 = e.z after the left-handed
// and odd-order negations of matrix3.cpp:183-197 cancel. Columns scaled by s, translation in the last column.
Mat4 TransformationMat4(const Vec3 &p, const Vec3 &e, const Vec3 &s) {
	const float s0 = sinf(e.y), c0 = cosf(e.y), s1 = sinf(e.x), c1 = cosf(e.x), s2 = sinf(e.z), c2 = cosf(e.z);
	const float s0s2 = s0 * s2, s0c2 = s0 * c2, c0s2 = c0 * s2, c0c2 = c0 * c2;
	const float r11 = c1 * c2, r10 = c1 * s2, r12 = -s1;
	const float r01 = s1 * s0c2 - c0s2, r00 = s1 * s0s2 + c0c2, r02 = s0 * c1;
	const float r21 = s1 * c0c2 + s0s2, r20 = s1 * c0s2 - s0c2, r22 = c0 * c1;
	Mat4 o;
	o.m[0][0] = r00 * s.x, o.m[0][1] = r01 * s.y, o.m[0][2] = r02 * s.z, o.m[0][3] = p.x;
	o.m[1][0] = r10 * s.x, o.m[1][1] = r11 * s.y, o.m[1][2] = r12 * s.z, o.m[1][3] = p.y;
	o.m[2][0] = r20 * s.x, o.m[2][1] = r21 * s.y, o.m[2][2] = r22 * s.z, o.m[2][3] = p.z;
	return o;
}

// ---- projection (projection.cpp:9-37). Mat44 here is m[row][col]. ----
static bool g_ndc_origin_bottom_left = false, g_ndc_homogeneous_depth = false;
void SetNDCInfos(bool origin_bottom_left, bool homogeneous_depth) { g_ndc_origin_bottom_left = origin_bottom_left, g_ndc_homogeneous_depth = homogeneous_depth; }

float FovToZoomFactor(float fov) {
	const float lo = 0.1f / 180.f * 3.1415926535f, hi = 179.9f / 180.f * 3.1415926535f;
	const float cf = fov < lo ? lo : (fov > hi ? hi : fov);
	return 1.f / float(tan(double(cf * 0.5f))); // math.cpp:93-95: Tan() is the DOUBLE tan, rounded
}

Mat44 ComputePerspectiveProjectionMatrix(float znear, float zfar, float zoom, const Vec2 &ar, const Vec2 &offset, const Vec2 &center_offset) {
	const float qA = g_ndc_homogeneous_depth ? ((zfar + znear) / (zfar - znear)) : (zfar / (zfar - znear));
	const float qB = g_ndc_homogeneous_depth ? (-2 * zfar * znear / (zfar - znear)) : (-qA * znear);
	Mat44 o{};
	o.m[0][0] = zoom / ar.x;
	o.m[1][1] = zoom / ar.y;
	o.m[2][2] = qA, o.m[2][3] = qB;
	o.m[3][2] = 1.f;
	o.m[0][2] = center_offset.x, o.m[1][2] = center_offset.y; // third column
	o.m[0][3] = offset.x, o.m[1][3] = offset.y;               // fourth column
	return o;
}

Mat44 ComputeOrthographicProjectionMatrix(float znear, float zfar, float size, const Vec2 &ar, const Vec2 &offset) {
	const float qA = g_ndc_homogeneous_depth ? (2.f / (zfar - znear)) : (1.f / (zfar - znear));
	const float qB = g_ndc_homogeneous_depth ? (-(zfar + znear) / (zfar - znear)) : (-qA * znear);
	Mat44 o{};
	o.m[0][0] = 2.f / size / ar.x;
	o.m[1][1] = 2.f / size / ar.y;
	o.m[2][2] = qA, o.m[2][3] = qB;
	o.m[3][3] = 1.f;
	o.m[0][3] = offset.x, o.m[1][3] = offset.y;
	return o;
}

// ---- frustum (frustum.cpp:12-28): planes = negated, normalised W±Y, W±X, W±Z ----
static Vec4 neg_normalized(float x, float y, float z, float w) {
	const float l = sqrtf(x * x + y * y + z * z);
	return Vec4(-(x / l), -(y / l), -(z / l), -(w / l));
}

Frustum MakeFrustum(const Mat44 &p) {
	Frustum f;
	const int row_of[3] = {1, 0, 2}; // Top/Bottom from Y, Left/Right from X, Near/Far from Z
	for (int k = 0; k < 3; ++k) {
		const float *R = p.m[row_of[k]], *W = p.m[3];
		f[k * 2 + 0] = neg_normalized(W[0] + R[0], W[1] + R[1], W[2] + R[2], W[3] + R[3]);
		f[k * 2 + 1] = neg_normalized(W[0] - R[0], W[1] - R[1], W[2] - R[2], W[3] - R[3]);
	}
	return f;
}

Frustum TransformFrustum(const Frustum &frustum, const Mat4 &m) { // plane' = transpose(inverse(m)) * plane
	const Mat4 inv = InverseFast(m);
	float I[4][4] = {{inv.m[0][0], inv.m[0][1], inv.m[0][2], inv.m[0][3]}, {inv.m[1][0], inv.m[1][1], inv.m[1][2], inv.m[1][3]},
		{inv.m[2][0], inv.m[2][1], inv.m[2][2], inv.m[2][3]}, {0.f, 0.f, 0.f, 1.f}};
	Frustum o;
	for (int k = 0; k < FP_Count; ++k) {
		const Vec4 &v = frustum[k];
		float r[4];
		for (int c = 0; c < 4; ++c)
			r[c] = v.x * I[0][c] + v.y * I[1][c] + v.z * I[2][c] + v.w * I[3][c];
		o[k] = Vec4(r[0], r[1], r[2], r[3]);
	}
	return o;
}

Frustum MakeFrustum(const Mat44 &projection, const Mat4 &world) { return TransformFrustum(MakeFrustum(projection), world); }

// ---- view states (render_pipeline.cpp:67-81): frustum, projection and view matrix of a camera world matrix ----
ViewState ComputeOrthographicViewState(const Mat4 &world, float size, float znear, float zfar, const Vec2 &aspect_ratio, const Vec2 &offset) {
	ViewState vs;
	vs.view = InverseFast(world);
	vs.proj = ComputeOrthographicProjectionMatrix(znear, zfar, size, aspect_ratio, offset);
	vs.frustum = MakeFrustum(vs.proj, world);
	return vs;
}

ViewState ComputePerspectiveViewState(const Mat4 &world, float fov, float znear, float zfar, const Vec2 &aspect_ratio, const Vec2 &offset, const Vec2 &center_offset) {
	ViewState vs;
	vs.view = InverseFast(world);
	vs.proj = ComputePerspectiveProjectionMatrix(znear, zfar, FovToZoomFactor(fov), aspect_ratio, offset, center_offset);
	vs.frustum = MakeFrustum(vs.proj, world);
	return vs;
}

} // namespace hg
