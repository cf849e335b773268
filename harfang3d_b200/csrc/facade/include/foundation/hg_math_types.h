// hg::Scene facade — math value types of the scene-update path, bit-compatible with the reference's foundation types
// (harfang/foundation/vector3.h:18, vector4.h, matrix4.h:25, matrix44.h, minmax.h:13-18, quaternion.h, frustum.h:14-16).
// Host-side functions here are EDIT-TIME helpers (building camera matrices and frusta); the per-frame arithmetic of
// Scene::Update / CullModelDisplayLists runs on the GPU behind include/hgb200.h and nowhere else.
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>

namespace hg {

using time_ns = int64_t;
inline time_ns time_from_sec_f(float v) { return time_ns(double(v) * 1e9); }
inline time_ns time_from_ms(int64_t ms) { return ms * 1000000; }
inline float time_to_sec_f(time_ns t) { return float(double(t) / 1e9); } // foundation/time.h:13

struct Vec2 {
	float x{0}, y{0};
	Vec2() = default;
	Vec2(float x_, float y_) : x(x_), y(y_) {}
};

struct Vec3 {
	float x{0}, y{0}, z{0};
	Vec3() = default;
	Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
	static const Vec3 Zero, One;
};
inline Vec3 operator+(const Vec3 &a, const Vec3 &b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(const Vec3 &a, const Vec3 &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(const Vec3 &a, float k) { return {a.x * k, a.y * k, a.z * k}; }
inline bool operator==(const Vec3 &a, const Vec3 &b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

struct Vec4 {
	float x{0}, y{0}, z{0}, w{1};
	Vec4() = default;
	Vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
};

struct Color { // foundation/color.h:17-32
	float r{0}, g{0}, b{0}, a{0};
	Color() = default;
	Color(float r_, float g_, float b_, float a_ = 1.f) : r(r_), g(g_), b(b_), a(a_) {}
};
inline bool operator==(const Color &x, const Color &y) { return x.r == y.r && x.g == y.g && x.b == y.b && x.a == y.a; }

struct Quaternion {
	float x{0}, y{0}, z{0}, w{1};
	Quaternion() = default;
	Quaternion(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
};

// 3x4 affine matrix, row-major, implicit last row 0 0 0 1 (foundation/matrix4.h:25)
struct Mat4 {
	float m[3][4];
	Mat4() { memcpy(m, Identity.m, sizeof(m)); }
	explicit Mat4(const float *p) { memcpy(m, p, sizeof(m)); }
	static const Mat4 Identity, Zero;
private:
	struct raw_tag {};
	Mat4(raw_tag, float d) {
		memset(m, 0, sizeof(m));
		m[0][0] = m[1][1] = m[2][2] = d;
	}
};
inline bool operator==(const Mat4 &a, const Mat4 &b) { return memcmp(a.m, b.m, sizeof(a.m)) == 0; }

struct Mat44 {
	float m[4][4];
};

struct MinMax {
	Vec3 mn, mx;
	MinMax() = default;
	MinMax(const Vec3 &a, const Vec3 &b) : mn(a), mx(b) {}
};
inline MinMax MinMaxFromPositionSize(const Vec3 &p, const Vec3 &s) { return {p - s * 0.5f, p + s * 0.5f}; }

enum FrustumPlane { FP_Top, FP_Bottom, FP_Left, FP_Right, FP_Near, FP_Far, FP_Count };
using Frustum = std::array<Vec4, FP_Count>;
enum Visibility : uint8_t { V_Outside, V_Inside, V_Clipped };

// ---- edit-time helpers (reference operation order, so a frustum built here has the same bits as the reference's) ----
Mat4 operator*(const Mat4 &a, const Mat4 &b);                              // foundation/matrix4.cpp:31-44
Mat4 InverseFast(const Mat4 &m);                                           // foundation/matrix4.cpp:293-303
Mat4 TransformationMat4(const Vec3 &p, const Vec3 &e, const Vec3 &s = Vec3(1, 1, 1)); // matrix4.cpp:176-183, RO_YXZ
Mat4 TranslationMat4(const Vec3 &p);
float FovToZoomFactor(float fov);                                          // foundation/projection.cpp:19
Mat44 ComputePerspectiveProjectionMatrix(float znear, float zfar, float zoom_factor, const Vec2 &aspect_ratio, const Vec2 &offset = Vec2(),
	const Vec2 &center_offset = Vec2());                                    // projection.cpp:29-37
Mat44 ComputeOrthographicProjectionMatrix(float znear, float zfar, float size, const Vec2 &aspect_ratio, const Vec2 &offset = Vec2()); // projection.cpp:21-27
Frustum MakeFrustum(const Mat44 &projection);                              // foundation/frustum.cpp:17-23
Frustum MakeFrustum(const Mat44 &projection, const Mat4 &world);           // frustum.cpp:25
Frustum TransformFrustum(const Frustum &frustum, const Mat4 &mtx);         // frustum.cpp:27-28
void SetNDCInfos(bool origin_bottom_left, bool homogeneous_depth);         // projection.cpp:9-16 (default false,false)

struct ViewState { // engine/render_pipeline.h:55-59
	Frustum frustum;
	Mat44 proj;
	Mat4 view;
};
ViewState ComputeOrthographicViewState(const Mat4 &world, float size, float znear, float zfar, const Vec2 &aspect_ratio, const Vec2 &offset = Vec2()); // render_pipeline.cpp:67-73
ViewState ComputePerspectiveViewState(const Mat4 &world, float fov, float znear, float zfar, const Vec2 &aspect_ratio, const Vec2 &offset = Vec2(),
	const Vec2 &center_offset = Vec2()); // render_pipeline.cpp:75-81

} // namespace hg
