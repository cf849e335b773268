/* TEST INFRASTRUCTURE ONLY — plain-C restatement ("port") of the reference's scene-update hot path.
 *
 * Parity status: PINNED. Every function here is checked bit-for-bit against the unmodified reference compiled into
 * oracle/_ref/libhg_ref.so (tests/test_oracle_port_vs_ref.py, run where _ref exists) and against golden vectors
 * generated from it and from the reference's own unit tests (tests/golden/, tests/test_oracle_golden.py).
 * hgp_sinf/hgp_cosf are checked against the libm the reference links for ALL 2^32 float inputs
 * (oracle/exhaustive_sincos.c: 0 mismatches on glibc 2.39 / FMA-capable x86-64).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/--impl reference legs may load this. The
 * product (harfang3d_b200/) never does.
 *
 * Conventions: Mat4 = float[12], row-major 3x4 == hg::Mat4::m (matrix4.h:25). MinMax = float[6] {mn, mx}.
 * Frustum = float[24] = 6 planes x {x,y,z,w}, order Top,Bottom,Left,Right,Near,Far (frustum.h:14).
 * All single-precision arithmetic is unfused and in the reference's source order (build with -ffp-contract=off). */
#ifndef HG_PORT_H
#define HG_PORT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

float hgp_sinf(float x); /* glibc 2.39 sysdeps/ieee754/flt-32/s_sinf.c (+sincosf.h), FMA build */
float hgp_cosf(float x);

void hgp_transformation_mat4(const float *trs9, float *out12);              /* matrix4.cpp:176-183 + matrix3.cpp:180-260 (RO_YXZ) */
void hgp_mat4_mul(const float *a, const float *b, float *out12);            /* matrix4.cpp:31-44 */
void hgp_inverse_fast(const float *m, float *out12);                        /* matrix4.cpp:293-303 */
void hgp_minmax_transform(const float *m12, const float *mm6, float *out6); /* minmax.cpp:226-244 */
int hgp_test_visibility(const float *frustum24, const float *mm6);          /* frustum.cpp:59-79 : 0 outside 1 inside 2 clipped */
void hgp_make_perspective_frustum(float znear, float zfar, float fov, float ar_x, float ar_y, const float *cam_world12,
	float *out24);                                                          /* projection.cpp:19,29-37 + frustum.cpp:12-28 */
void hgp_quat_to_euler(const float *q4, float *out3);                       /* quaternion.cpp:57-60,141-168 + matrix3.cpp:91-163 */
void hgp_to_bgfx(const float *m12, float *out16);                           /* render_pipeline.cpp:59-62 */

/* animation.h:109-128 — returns 1 when t lies inside [keys[kf0], keys[kf1]) */
int hgp_get_interval_keys(int n_keys, const int64_t *t, int64_t at, int *kf0, int *kf1);
/* animation.h:155-169 + math.h:137-149; returns 0 (v untouched) when the track is empty */
int hgp_eval_vec3_track(int n_keys, const int64_t *t, const float *v3, const float *tension, const float *bias, int64_t at, float *out3);
/* animation.h:141-153 + math.h:114-116 on Quaternion (component-wise lerp) */
int hgp_eval_quat_track(int n_keys, const int64_t *t, const float *v4, int64_t at, float *out4);

/* scene.cpp:2380-2431 clock step of one playing anim. loop_mode: 0 once 1 infinite 2 loop. Returns 1 if the anim
 * must be removed after this frame's evaluation (ALM_Once reached its end). */
int hgp_step_play_clock(int64_t *t, int64_t dt, int8_t t_scale, int loop_mode, int64_t t_start, int64_t t_end, int paused);

/* scene.cpp:263-289 — worlds for n transform slots. parent[i] = transform slot of the parent or -1.
 * used[i]==0 -> slot skipped. preset[i]!=0 -> worlds[i] already holds a SetNodeWorldMatrix value and is kept. */
void hgp_compute_world_matrices(int n, const float *trs9, const int32_t *parent, const uint8_t *used, const uint8_t *preset, float *worlds12);

/* render_pipeline.cpp:2098-2104 on a flat entry table. bounds6 is indexed by bounds_idx[e]. Writes the surviving
 * entry indices in order; returns their count. */
int hgp_cull(int n_entries, const uint32_t *mtx_idx, const uint32_t *bounds_idx, const float *bounds6, const float *worlds12,
	const float *frustum24, uint32_t *out_visible);

/* render_pipeline.cpp:2172-2173 */
void hgp_bone_matrices(int n_bones, const uint32_t *mtx_idx, const uint32_t *bind_idx, const float *worlds12, const float *bind_pose12, float *out16);

#ifdef __cplusplus
}
#endif
#endif
