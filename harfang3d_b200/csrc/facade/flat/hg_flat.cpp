// Flat C binding over the hg::Scene API: builds a scene from plain arrays, runs frames, reads worlds / display lists / visible lists /
// bone matrices back — what a ctypes / FFI caller (tests, bench.py) needs instead of C++ handles. The file uses nothing but the PUBLIC
// hg::Scene interface, so it compiles unchanged against two implementations of it:
//   * the product's hg::Scene facade (csrc/facade/include, -DHG_FACADE)  -> harfang3d_b200/libhg_facade_flat.so (`make facade_flat`);
//     this is the library bench.py's `e2e_facade` leg and tests/test_facade.py drive;
//   * the UNMODIFIED reference headers + sources under /root/reference    -> the oracle's build of the reference (`make -C oracle ref`), the parity
//     oracle and the CPU baseline: the reference's own
//       Scene::Update                (harfang/engine/scene.cpp:307-315)
//       Scene::GetModelDisplayLists  (harfang/engine/scene.cpp:318-404)
//       CullModelDisplayLists        (harfang/engine/render_pipeline.cpp:2098-2104)
//     behind the very calls that drive the facade. The only restated arithmetic in that build is the bone-matrix product
//     (render_pipeline.cpp:2172-2173 + to_bgfx :59-62): the reference function is static and interleaved with bgfx calls.
// Entry points keep the hgref_ prefix in both builds (one ctypes table); the probes of reference-internal host math are compiled out
// of the facade build, the fused-frame entry points (hgfacade_*) out of the reference build.
#include "engine/animation.h"
#include "engine/render_pipeline.h"
#include "engine/scene.h"
#include "foundation/frustum.h"
#include "foundation/matrix3.h"
#include "foundation/matrix4.h"
#include "foundation/matrix44.h"
#include "foundation/minmax.h"
#include "foundation/projection.h"
#include "foundation/quaternion.h"
#include "foundation/log.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <limits>
#include <vector>

using namespace hg;

namespace {

struct RefCtx {
	PipelineResources resources;
	std::unique_ptr<Scene> scene;
	std::vector<ModelRef> models;
	std::vector<NodeRef> nodes; // by creation order == flat node index
	std::vector<AnimRef> anims;
	std::vector<SceneAnimRef> scene_anims;
	std::vector<ScenePlayAnimRef> play_anims;

	std::vector<ModelDisplayList> opaque, transparent, culled;
	std::vector<SkinnedModelDisplayList> opaque_skinned, transparent_skinned;

	RefCtx() : scene(new Scene) {}
	~RefCtx() { scene.reset(); }
};

inline Mat4 mat4_from(const float *m) { // 12 floats, row-major 3x4 == hg::Mat4::m
	Mat4 o;
	memcpy(o.m, m, sizeof(float) * 12);
	return o;
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

} // namespace

#ifdef HG_FACADE
#include "hgb200.h"
namespace hg {
extern double g_visible_breakdown[3];
}
#endif
extern "C" {

void *hgref_create() {
	set_log_level(LL_Error); // the path warns on invalid refs (scene.cpp:750-766); keep stderr quiet in tests
	return new RefCtx;
}

void hgref_destroy(void *h) { delete static_cast<RefCtx *>(h); }

// Model fixture exactly like harfang/tests/engine/scene.cpp:116-122: bounds + lists of BGFX_INVALID_HANDLEs + mats.
int hgref_add_model(void *h, int n_lists, const float *bounds6, const uint16_t *mats, const int *bone_off, const uint16_t *bone_table, int n_bind,
	const float *bind_pose12) {
	auto *c = static_cast<RefCtx *>(h);
	Model mdl;
	for (int i = 0; i < n_lists; ++i) {
		mdl.bounds.push_back(MinMax(Vec3(bounds6[i * 6 + 0], bounds6[i * 6 + 1], bounds6[i * 6 + 2]), Vec3(bounds6[i * 6 + 3], bounds6[i * 6 + 4], bounds6[i * 6 + 5])));
		DisplayList dl{BGFX_INVALID_HANDLE, BGFX_INVALID_HANDLE, {}};
		if (bone_off)
			for (int j = bone_off[i]; j < bone_off[i + 1]; ++j)
				dl.bones_table.push_back(bone_table[j]);
		mdl.lists.push_back(dl);
		mdl.mats.push_back(mats[i]);
	}
	for (int i = 0; i < n_bind; ++i)
		mdl.bind_pose.push_back(mat4_from(bind_pose12 + i * 12));
	char name[32];
	snprintf(name, sizeof(name), "mdl%d", int(c->models.size()));
	c->models.push_back(c->resources.models.Add(name, mdl));
	return int(c->models.size()) - 1;
}

// Append n nodes. parent[] = flat node index (may be > own index) or -1; resolved in a second pass.
// model[] = -1 -> no object component. has_transform may be NULL (all true).
int hgref_add_nodes(void *h, int n, const float *trs9, const int32_t *parent, const uint8_t *disabled, const int32_t *model, const uint8_t *n_mats,
	const uint32_t *transparent_mask, const uint8_t *has_transform) {
	auto *c = static_cast<RefCtx *>(h);
	auto &scene = *c->scene;
	const size_t base = c->nodes.size();
	scene.ReserveNodes(base + n);
	scene.ReserveTransforms(base + n);
	for (int i = 0; i < n; ++i) {
		Node node = scene.CreateNode();
		if (!has_transform || has_transform[i]) {
			const float *t = trs9 + size_t(i) * 9;
			node.SetTransform(scene.CreateTransform(Vec3(t[0], t[1], t[2]), Vec3(t[3], t[4], t[5]), Vec3(t[6], t[7], t[8])));
		}
		if (model && model[i] >= 0) {
			std::vector<Material> mats(n_mats ? n_mats[i] : 1);
			for (size_t k = 0; k < mats.size(); ++k)
				if (transparent_mask && (transparent_mask[i] >> k) & 1)
					SetMaterialBlendMode(mats[k], BM_Alpha);
			node.SetObject(scene.CreateObject(c->models[model[i]], mats));
		}
		if (disabled && disabled[i])
			node.Disable();
		c->nodes.push_back(node.ref);
	}
	for (int i = 0; i < n; ++i)
		if (parent && parent[i] >= 0) {
			const auto trs_ref = scene.GetNodeTransformRef(c->nodes[base + i]);
			if (scene.IsValidTransformRef(trs_ref))
				scene.SetTransformParent(trs_ref, c->nodes[parent[i]]);
		}
	return int(c->nodes.size());
}

int hgref_set_object_bones(void *h, int node, int n, const int32_t *bone_nodes) {
	auto *c = static_cast<RefCtx *>(h);
	const auto obj = c->scene->GetNodeObjectRef(c->nodes[node]);
	c->scene->SetObjectBoneCount(obj, n);
	for (int i = 0; i < n; ++i)
		c->scene->SetObjectBone(obj, i, bone_nodes[i] >= 0 ? c->nodes[bone_nodes[i]] : InvalidNodeRef);
	return 0;
}

void hgref_set_trs(void *h, int node, const float *t) {
	auto *c = static_cast<RefCtx *>(h);
	c->scene->SetTransformTRS(c->scene->GetNodeTransformRef(c->nodes[node]), {Vec3(t[0], t[1], t[2]), Vec3(t[3], t[4], t[5]), Vec3(t[6], t[7], t[8])});
}

void hgref_get_trs(void *h, int node, float *t) {
	auto *c = static_cast<RefCtx *>(h);
	const auto trs = c->scene->GetTransformTRS(c->scene->GetNodeTransformRef(c->nodes[node]));
	memcpy(t, &trs.pos, 12);
	memcpy(t + 3, &trs.rot, 12);
	memcpy(t + 6, &trs.scl, 12);
}

void hgref_set_parent(void *h, int node, int parent) {
	auto *c = static_cast<RefCtx *>(h);
	c->scene->SetTransformParent(c->scene->GetNodeTransformRef(c->nodes[node]), parent >= 0 ? c->nodes[parent] : InvalidNodeRef);
}

void hgref_enable_node(void *h, int node, int enable) {
	auto *c = static_cast<RefCtx *>(h);
	enable ? c->scene->EnableNode(c->nodes[node]) : c->scene->DisableNode(c->nodes[node]);
}

void hgref_destroy_node(void *h, int node) { static_cast<RefCtx *>(h)->scene->DestroyNode(static_cast<RefCtx *>(h)->nodes[node]); }
int hgref_garbage_collect(void *h) { return int(static_cast<RefCtx *>(h)->scene->GarbageCollect()); }

// scene.cpp:768-786 — pre-set world matrix (wins over ComputeWorldMatrices when called after ReadyWorldMatrices)
void hgref_set_node_world(void *h, int node, const float *m12) {
	auto *c = static_cast<RefCtx *>(h);
	c->scene->SetNodeWorldMatrix(c->nodes[node], mat4_from(m12));
}

int hgref_node_transform_idx(void *h, int node) {
	auto *c = static_cast<RefCtx *>(h);
	const auto r = c->scene->GetNodeTransformRef(c->nodes[node]);
	return c->scene->IsValidTransformRef(r) ? int(r.idx) : -1;
}

// ---- animation -----------------------------------------------------------------------------------------------
// One hg::Anim per animated node. Track layout of the flat description (shared with the product's C ABI):
//   vec3 Hermite key  = { int64 t; float v[3]; float tension; float bias; }  (animation.h:42-46)  -> 28 B packed as arrays
//   quat linear key   = { int64 t; float v[4]; }                             (animation.h:28-31)
// kinds: 0 Position, 1 Rotation (vec3 Euler), 2 Scale, 3 Rotation (quaternion, sets AF_UseQuaternionForRotation)
int hgref_add_anim(void *h, int use_quat) {
	auto *c = static_cast<RefCtx *>(h);
	Anim anim;
	anim.flags = use_quat ? AF_UseQuaternionForRotation : 0;
	c->anims.push_back(c->scene->AddAnim(anim));
	return int(c->anims.size()) - 1;
}

int hgref_anim_add_vec3_track(void *h, int anim, int kind, int n_keys, const int64_t *t, const float *v3, const float *tension, const float *bias) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	AnimTrackHermiteT<Vec3> trk;
	trk.target = kind == 0 ? "Position" : (kind == 1 ? "Rotation" : "Scale");
	for (int i = 0; i < n_keys; ++i) {
		AnimKeyHermiteT<Vec3> k;
		k.t = t[i];
		k.v = Vec3(v3[i * 3], v3[i * 3 + 1], v3[i * 3 + 2]);
		k.tension = tension ? tension[i] : 0.f;
		k.bias = bias ? bias[i] : 0.f;
		trk.keys.push_back(k);
	}
	a->vec3_tracks.push_back(trk);
	return int(a->vec3_tracks.size()) - 1;
}

// float Hermite track with a free target name ("Camera.Fov": BindNodeAnim scene.cpp:2091-2101, evaluated by EvaluateBoundAnim :2194-2197)
int hgref_anim_add_float_track(void *h, int anim, const char *target, int n_keys, const int64_t *t, const float *v, const float *tension, const float *bias) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	AnimTrackHermiteT<float> trk;
	trk.target = target;
	for (int i = 0; i < n_keys; ++i) {
		AnimKeyHermiteT<float> k;
		k.t = t[i];
		k.v = v[i];
		k.tension = tension ? tension[i] : 0.f;
		k.bias = bias ? bias[i] : 0.f;
		trk.keys.push_back(k);
	}
	a->float_tracks.push_back(trk);
	return int(a->float_tracks.size()) - 1;
}

// "Enable" bool track (BindNodeAnim scene.cpp:2083-2089, evaluated by EvaluateBoundAnim :2158-2162 with EvaluateStep)
int hgref_anim_add_enable_track(void *h, int anim, int n_keys, const int64_t *t, const uint8_t *v) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	AnimTrackT<bool> trk;
	trk.target = "Enable";
	for (int i = 0; i < n_keys; ++i)
		trk.keys.push_back({t[i], v[i] != 0});
	a->bool_tracks.push_back(trk);
	return int(a->bool_tracks.size()) - 1;
}
int hgref_is_node_enabled(void *h, int node) { return static_cast<RefCtx *>(h)->scene->IsNodeEnabled(static_cast<RefCtx *>(h)->nodes[node]) ? 1 : 0; }

int hgref_anim_add_quat_track(void *h, int anim, int n_keys, const int64_t *t, const float *v4) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	AnimTrackT<Quaternion> trk;
	trk.target = "Rotation";
	for (int i = 0; i < n_keys; ++i)
		trk.keys.push_back({t[i], Quaternion(v4[i * 4], v4[i * 4 + 1], v4[i * 4 + 2], v4[i * 4 + 3])});
	a->quat_tracks.push_back(trk);
	return int(a->quat_tracks.size()) - 1;
}

// Bulk form of the three calls above (one ctypes call for a whole flat description; used for the large bench scenes).
int hgref_add_anims_bulk(void *h, int n_anims, const uint8_t *use_quat, int n_tracks, const int32_t *trk_anim, const uint8_t *trk_kind, const uint32_t *key_off,
	const uint32_t *key_count, const int64_t *v3_t, const float *v3_val, const float *v3_tension, const float *v3_bias, const int64_t *q_t, const float *q_val) {
	auto *c = static_cast<RefCtx *>(h);
	const int base = int(c->anims.size());
	for (int a = 0; a < n_anims; ++a)
		hgref_add_anim(h, use_quat ? use_quat[a] : 0);
	for (int k = 0; k < n_tracks; ++k) {
		const size_t o = key_off[k];
		if (trk_kind[k] == 3)
			hgref_anim_add_quat_track(h, base + trk_anim[k], int(key_count[k]), q_t + o, q_val + o * 4);
		else
			hgref_anim_add_vec3_track(h, base + trk_anim[k], trk_kind[k], int(key_count[k]), v3_t + o, v3_val + o * 3, v3_tension ? v3_tension + o : nullptr,
				v3_bias ? v3_bias + o : nullptr);
	}
	return int(c->anims.size());
}

// SceneAnim = list of (node, anim) pairs (scene.h:104-114)
int hgref_add_scene_anim(void *h, int n, const int32_t *node, const int32_t *anim, int64_t t_start, int64_t t_end) {
	auto *c = static_cast<RefCtx *>(h);
	SceneAnim sa;
	sa.t_start = t_start;
	sa.t_end = t_end;
	for (int i = 0; i < n; ++i)
		sa.node_anims.push_back({c->nodes[node[i]], c->anims[anim[i]]});
	c->scene_anims.push_back(c->scene->AddSceneAnim(sa));
	return int(c->scene_anims.size()) - 1;
}

// Scene::PlayAnim scene.cpp:2353-2375; loop_mode: 0 once, 1 infinite, 2 loop (animation.h:26)
int hgref_play_anim(void *h, int scene_anim, int loop_mode, float t_scale) {
	auto *c = static_cast<RefCtx *>(h);
	c->play_anims.push_back(c->scene->PlayAnim(c->scene_anims[scene_anim], AnimLoopMode(loop_mode), E_Linear, UnspecifiedAnimTime, UnspecifiedAnimTime, false, t_scale));
	return int(c->play_anims.size()) - 1;
}

int hgref_is_playing(void *h, int play) {
	auto *c = static_cast<RefCtx *>(h);
	return c->scene->IsPlaying(c->play_anims[play]) ? 1 : 0;
}

// ---- lights, material values, scene instances (single-value animation targets and NF_InstanceDisabled propagation) --------------------
int hgref_add_point_light(void *h, int node, float radius) {
	auto *c = static_cast<RefCtx *>(h);
	c->scene->SetNodeLight(c->nodes[node], c->scene->CreatePointLight(radius).ref);
	return 1;
}
// out10: diffuse rgba, diffuse intensity, specular rgba, specular intensity
void hgref_get_light(void *h, int node, float *out10) {
	auto *c = static_cast<RefCtx *>(h);
	const auto l = c->scene->GetNodeLightRef(c->nodes[node]);
	const Color d = c->scene->GetLightDiffuseColor(l), s = c->scene->GetLightSpecularColor(l);
	const float o[10] = {d.r, d.g, d.b, d.a, c->scene->GetLightDiffuseIntensity(l), s.r, s.g, s.b, s.a, c->scene->GetLightSpecularIntensity(l)};
	memcpy(out10, o, sizeof(o));
}
int hgref_anim_add_color_track(void *h, int anim, const char *target, int n_keys, const int64_t *t, const float *v4, const float *tension, const float *bias) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	AnimTrackHermiteT<Color> trk;
	trk.target = target;
	for (int i = 0; i < n_keys; ++i) {
		AnimKeyHermiteT<Color> k;
		k.t = t[i];
		k.v = Color(v4[i * 4], v4[i * 4 + 1], v4[i * 4 + 2], v4[i * 4 + 3]);
		k.tension = tension ? tension[i] : 0.f;
		k.bias = bias ? bias[i] : 0.f;
		trk.keys.push_back(k);
	}
	a->color_tracks.push_back(trk);
	return int(a->color_tracks.size()) - 1;
}
int hgref_anim_add_vec4_track(void *h, int anim, const char *target, int n_keys, const int64_t *t, const float *v4, const float *tension, const float *bias) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	AnimTrackHermiteT<Vec4> trk;
	trk.target = target;
	for (int i = 0; i < n_keys; ++i) {
		AnimKeyHermiteT<Vec4> k;
		k.t = t[i];
		k.v = Vec4(v4[i * 4], v4[i * 4 + 1], v4[i * 4 + 2], v4[i * 4 + 3]);
		k.tension = tension ? tension[i] : 0.f;
		k.bias = bias ? bias[i] : 0.f;
		trk.keys.push_back(k);
	}
	a->vec4_tracks.push_back(trk);
	return int(a->vec4_tracks.size()) - 1;
}
void hgref_set_material_value(void *h, int node, int slot, const char *name, const float *v4) {
	auto *c = static_cast<RefCtx *>(h);
	// written straight into the value map: the reference's SetMaterialValue also creates a bgfx uniform, which needs a renderer
	auto &mv = c->scene->GetObjectMaterial(c->scene->GetNodeObjectRef(c->nodes[node]), size_t(slot)).values[name];
	mv.value = {v4[0], v4[1], v4[2], v4[3]};
}
int hgref_get_material_value(void *h, int node, int slot, const char *name, float *out4) {
	auto *c = static_cast<RefCtx *>(h);
	Material &m = c->scene->GetObjectMaterial(c->scene->GetNodeObjectRef(c->nodes[node]), size_t(slot));
	const auto i = m.values.find(name);
	if (i == m.values.end() || i->second.value.size() < 4)
		return 0;
	memcpy(out4, i->second.value.data(), 16);
	return 1;
}
void hgref_set_scene_anim_name(void *h, int scene_anim, const char *name) {
	auto *c = static_cast<RefCtx *>(h);
	if (SceneAnim *sa = c->scene->GetSceneAnim(c->scene_anims[scene_anim]))
		sa->name = name;
}
// an Instance component on `node`: scene file `path`, anim started on instantiation (may be empty)
int hgref_add_instance(void *h, int node, const char *path, const char *anim, int loop_mode) {
	auto *c = static_cast<RefCtx *>(h);
	Instance inst = c->scene->CreateInstance(path);
	inst.SetOnInstantiateAnim(anim ? anim : "");
	inst.SetOnInstantiateAnimLoopMode(AnimLoopMode(loop_mode));
	c->scene->SetNodeInstance(c->nodes[node], inst.ref);
	return inst.IsValid() ? 1 : 0;
}
// NodeSetupInstanceFromFile (scene.cpp:1304-1348,1390) + NodeStartOnInstantiateAnim (:1350-1373). The instantiated nodes / anims / scene
// anims are appended to the context's flat indices, in view order. Returns the number of instantiated nodes, -1 on failure.
int hgref_setup_instance(void *h, int node) {
	auto *c = static_cast<RefCtx *>(h);
	PipelineInfo pipeline;
	if (!c->scene->NodeSetupInstanceFromFile(c->nodes[node], c->resources, pipeline, LSSF_All | LSSF_DoNotLoadResources, 0))
		return -1;
	const SceneView &view = c->scene->GetNodeInstanceSceneView(c->nodes[node]);
	for (const auto &r : view.nodes)
		c->nodes.push_back(r);
	for (const auto &r : view.anims)
		c->anims.push_back(r);
	for (const auto &r : view.scene_anims)
		c->scene_anims.push_back(r);
	const auto play = c->scene->NodeStartOnInstantiateAnim(c->nodes[node]);
	if (c->scene->IsPlaying(play))
		c->play_anims.push_back(play);
	return int(view.nodes.size());
}
int hgref_node_count(void *h) { return int(static_cast<RefCtx *>(h)->nodes.size()); }
// instance_anim_track of an Anim: names = n_keys zero-terminated strings back to back
int hgref_anim_set_instance_track(void *h, int anim, int n_keys, const int64_t *t, const char *names, const uint8_t *loop_mode, const float *t_scale) {
	auto *c = static_cast<RefCtx *>(h);
	Anim *a = c->scene->GetAnim(c->anims[anim]);
	a->instance_anim_track.keys.clear();
	for (int i = 0; i < n_keys; ++i) {
		AnimKeyT<InstanceAnimKey> k;
		k.t = t[i];
		k.v.anim_name = names;
		names += strlen(names) + 1;
		k.v.loop_mode = AnimLoopMode(loop_mode[i]);
		k.v.t_scale = t_scale[i];
		a->instance_anim_track.keys.push_back(k);
	}
	return n_keys;
}

// ---- per-frame path ------------------------------------------------------------------------------------------
void hgref_update(void *h, int64_t dt) { static_cast<RefCtx *>(h)->scene->Update(dt); }

// The split form used by SceneUpdateSystemsImpl (scene_systems.cpp:153-179): phase 0 = Store+Ready+Anims, phase 1 = Compute+Fixup
void hgref_update_phase(void *h, int phase, int64_t dt) {
	auto &s = *static_cast<RefCtx *>(h)->scene;
	if (phase == 0) {
		s.StorePreviousWorldMatrices();
		s.ReadyWorldMatrices();
		s.UpdatePlayingAnims(dt);
	} else {
		s.ComputeWorldMatrices();
		s.FixupPreviousWorldMatrices();
	}
}

int hgref_world_count(void *h) { return int(static_cast<RefCtx *>(h)->scene->GetTransformWorldMatrices().size()); }

int hgref_get_worlds(void *h, float *out12, int previous) {
	auto &s = *static_cast<RefCtx *>(h)->scene;
	const auto &w = previous ? s.GetPreviousTransformWorldMatrices() : s.GetTransformWorldMatrices();
	if (out12 && !w.empty())
		memcpy(out12, w.data(), w.size() * sizeof(Mat4));
	return int(w.size());
}

// kind: 0 opaque, 1 transparent, 2 opaque skinned, 3 transparent skinned. Returns the 4 counts.
void hgref_get_model_display_lists(void *h, int *counts4) {
	auto *c = static_cast<RefCtx *>(h);
	c->scene->GetModelDisplayLists(c->opaque, c->transparent, c->opaque_skinned, c->transparent_skinned, c->resources);
	counts4[0] = int(c->opaque.size());
	counts4[1] = int(c->transparent.size());
	counts4[2] = int(c->opaque_skinned.size());
	counts4[3] = int(c->transparent_skinned.size());
}

static int dump_lists(const std::vector<ModelDisplayList> &v, uint32_t *mtx_idx, uint16_t *mdl_idx, uint16_t *lst_idx) {
	for (size_t i = 0; i < v.size(); ++i) {
		if (mtx_idx) mtx_idx[i] = v[i].mtx_idx;
		if (mdl_idx) mdl_idx[i] = v[i].mdl_idx;
		if (lst_idx) lst_idx[i] = v[i].lst_idx;
	}
	return int(v.size());
}

int hgref_read_display_lists(void *h, int kind, uint32_t *mtx_idx, uint16_t *mdl_idx, uint16_t *lst_idx) {
	auto *c = static_cast<RefCtx *>(h);
	return dump_lists(kind == 0 ? c->opaque : c->transparent, mtx_idx, mdl_idx, lst_idx);
}

// skinned: per list 32 mtx_idxs + 32 bones_idxs + bone_count/mdl/lst (render_pipeline.h:596-603)
int hgref_read_skinned_lists(void *h, int kind, uint32_t *mtx_idxs32, uint32_t *bones_idxs32, uint16_t *bone_count, uint16_t *mdl_idx, uint16_t *lst_idx) {
	auto *c = static_cast<RefCtx *>(h);
	const auto &v = kind == 2 ? c->opaque_skinned : c->transparent_skinned;
	for (size_t i = 0; i < v.size(); ++i) {
		memcpy(mtx_idxs32 + i * 32, v[i].mtx_idxs, 32 * 4);
		memcpy(bones_idxs32 + i * 32, v[i].bones_idxs, 32 * 4);
		bone_count[i] = v[i].bone_count;
		mdl_idx[i] = v[i].mdl_idx;
		lst_idx[i] = v[i].lst_idx;
	}
	return int(v.size());
}

// copy + CullModelDisplayLists exactly as scene_forward_pipeline.cpp:555-558 does; returns survivors, and for each survivor
// its position in the un-culled list (src_pos) — the "visible index list" of the metric.
int hgref_cull(void *h, int kind, const float *frustum24, uint32_t *src_pos, uint32_t *mtx_idx) {
	auto *c = static_cast<RefCtx *>(h);
	const auto &all = kind == 0 ? c->opaque : c->transparent;
	Frustum f;
	memcpy(f.data(), frustum24, 96);
	c->culled = all;
	CullModelDisplayLists(f, c->culled, c->scene->GetTransformWorldMatrices(), c->resources);
	// stable compaction => survivors are a subsequence of `all`; recover positions by a merge walk
	size_t j = 0;
	for (size_t i = 0; i < c->culled.size(); ++i) {
		while (memcmp(&all[j], &c->culled[i], sizeof(ModelDisplayList)) != 0)
			++j;
		if (src_pos) src_pos[i] = uint32_t(j);
		if (mtx_idx) mtx_idx[i] = c->culled[i].mtx_idx;
		++j;
	}
	return int(c->culled.size());
}

// restated: render_pipeline.cpp:2172-2173 (+ :2177-2181 for previous) and to_bgfx :59-62, using the reference's own operator*
int hgref_bone_matrices(void *h, int kind, int previous, float *out16) {
	auto *c = static_cast<RefCtx *>(h);
	const auto &v = kind == 2 ? c->opaque_skinned : c->transparent_skinned;
#ifdef HG_FACADE
	{ // the facade exposes the loop as a callable (computed on the device)
		std::vector<float> o;
		const size_t nb = ComputeSkinMatrices(v, previous ? c->scene->GetPreviousTransformWorldMatrices() : c->scene->GetTransformWorldMatrices(), c->resources,
			o, previous != 0);
		if (out16 && nb)
			memcpy(out16, o.data(), nb * 64);
		return int(nb);
	}
#else
	const auto &mtxs = previous ? c->scene->GetPreviousTransformWorldMatrices() : c->scene->GetTransformWorldMatrices();
	size_t n = 0;
	for (const auto &dl : v) {
		const auto &mdl = c->resources.models.Get_unsafe_(dl.mdl_idx);
		for (int j = 0; j < dl.bone_count; ++j, ++n) {
			const Mat4 m = mtxs[dl.mtx_idxs[j]] * mdl.bind_pose[dl.bones_idxs[j]];
			const float o[16] = {m.m[0][0], m.m[1][0], m.m[2][0], 0.f, m.m[0][1], m.m[1][1], m.m[2][1], 0.f, m.m[0][2], m.m[1][2], m.m[2][2], 0.f, m.m[0][3],
				m.m[1][3], m.m[2][3], 1.f};
			if (out16) memcpy(out16 + n * 16, o, 64);
		}
	}
	return int(n);
#endif
}

// ---- view / frustum (a12.3) ------------------------------------------------------------------------------------
// ComputePerspectiveProjectionMatrix (projection.cpp:29-37) + MakeFrustum(proj, world) (frustum.cpp:17-28).
// ComputePerspectiveViewState itself is avoided because it calls bgfx::getCaps() (render_pipeline.cpp:76).
void hgref_make_perspective_frustum(float znear, float zfar, float fov, float ar_x, float ar_y, const float *cam_world12, float *out24) {
	const Mat44 proj = ComputePerspectiveProjectionMatrix(znear, zfar, FovToZoomFactor(fov), Vec2(ar_x, ar_y));
	const Frustum f = MakeFrustum(proj, mat4_from(cam_world12));
	memcpy(out24, f.data(), 96);
}

void hgref_make_ortho_frustum(float znear, float zfar, float size, float ar_x, float ar_y, const float *cam_world12, float *out24) {
	const Mat44 proj = ComputeOrthographicProjectionMatrix(znear, zfar, size, Vec2(ar_x, ar_y));
	const Frustum f = MakeFrustum(proj, mat4_from(cam_world12));
	memcpy(out24, f.data(), 96);
}

// ---- camera component + view state (scene.h:295-324,594-599; scene.cpp:240-263; render_pipeline.cpp:67-81) ----
// kind 0: perspective (p = fov), 1: orthographic (p = size)
int hgref_set_node_camera(void *h, int node, int kind, float znear, float zfar, float p) {
	auto *c = static_cast<RefCtx *>(h);
	Camera cam = kind == 1 ? c->scene->CreateOrthographicCamera(znear, zfar, p) : c->scene->CreateCamera(znear, zfar, p);
	c->scene->SetNodeCamera(c->nodes[node], cam.ref);
	return cam.IsValid() ? 1 : 0;
}
void hgref_set_camera_fov(void *h, int node, float fov) {
	auto *c = static_cast<RefCtx *>(h);
	c->scene->SetCameraFov(c->scene->GetNodeCameraRef(c->nodes[node]), fov);
}
float hgref_get_camera_fov(void *h, int node) {
	auto *c = static_cast<RefCtx *>(h);
	return c->scene->GetCameraFov(c->scene->GetNodeCameraRef(c->nodes[node]));
}
void hgref_set_current_camera(void *h, int node) { static_cast<RefCtx *>(h)->scene->SetCurrentCamera(static_cast<RefCtx *>(h)->nodes[node]); }
// out: 24 frustum plane floats, 16 projection floats (m[row][col]), 12 view floats
void hgref_current_camera_view_state(void *h, float ar_x, float ar_y, float *out52) {
	const ViewState vs = static_cast<RefCtx *>(h)->scene->ComputeCurrentCameraViewState(Vec2(ar_x, ar_y));
	memcpy(out52, vs.frustum.data(), 96);
	memcpy(out52 + 24, vs.proj.m, 64);
	memcpy(out52 + 40, vs.view.m, 48);
}

// ---- math probes: let tests pin the port / GPU path against the reference's own functions and golden vectors ----
void hgref_transformation_mat4(const float *trs9, float *out12) {
	const Mat4 m = TransformationMat4(Vec3(trs9[0], trs9[1], trs9[2]), Vec3(trs9[3], trs9[4], trs9[5]), Vec3(trs9[6], trs9[7], trs9[8]));
	memcpy(out12, m.m, 48);
}
void hgref_mat4_mul(const float *a12, const float *b12, float *out12) {
	const Mat4 m = mat4_from(a12) * mat4_from(b12);
	memcpy(out12, m.m, 48);
}
void hgref_inverse_fast(const float *a12, float *out12) {
	const Mat4 m = InverseFast(mat4_from(a12));
	memcpy(out12, m.m, 48);
}
// Scene binary round trip. hgref_save_binary (reference build only): Scene::Save_binary through SaveSceneBinaryToFile (engine/scene.cpp:1729-1736,
// engine/scene_load_binary.cpp:237-541). hgref_load_binary: LoadSceneBinaryFromFile (scene.cpp:1803-1811 -> Load_binary, scene_load_binary.cpp:549-905)
// into this context's scene, with LSSF_DoNotLoadResources so that model names resolve to the models already registered through hgref_add_model
// (ResourceCache::Add returns the existing ref, engine/resource_cache.h:28-36). Nodes / anims / scene anims of the file become this
// context's flat indices, in file order. Returns the number of nodes loaded, -1 on failure.
#ifndef HG_FACADE
int hgref_save_binary(void *h, const char *path) {
	auto *c = static_cast<RefCtx *>(h);
	return SaveSceneBinaryToFile(path, *c->scene, c->resources, LSSF_All) ? 1 : 0;
}
#endif
int hgref_load_binary(void *h, const char *path) {
	auto *c = static_cast<RefCtx *>(h);
	LoadSceneContext lctx;
	PipelineInfo pipeline;
	if (!LoadSceneBinaryFromFile(path, *c->scene, c->resources, pipeline, lctx, LSSF_All | LSSF_DoNotLoadResources))
		return -1;
	for (const auto &r : lctx.view.nodes)
		c->nodes.push_back(r);
	for (const auto &r : lctx.view.anims)
		c->anims.push_back(r);
	for (const auto &r : lctx.view.scene_anims)
		c->scene_anims.push_back(r);
	return int(lctx.view.nodes.size());
}
// Scene::GetMinMax (engine/scene.cpp:1982) when count < 0, else GetNodesMinMax (engine/node.cpp:265-285) over nodes [first, first+count).
// Returns 1 and the box {mn.xyz, mx.xyz} when at least one node contributed.
int hgref_minmax(void *h, int first, int count, float *out6) {
	auto *c = static_cast<RefCtx *>(h);
	MinMax mm;
	bool ok;
	if (count < 0) {
		ok = c->scene->GetMinMax(c->resources, mm);
	} else {
		std::vector<Node> nodes;
		for (int i = first; i < first + count; ++i)
			nodes.push_back(c->scene->GetNode(c->nodes[i]));
		ok = GetNodesMinMax(nodes, c->resources, mm);
	}
	if (ok) {
		memcpy(out6, &mm.mn, 12);
		memcpy(out6 + 3, &mm.mx, 12);
	}
	return ok ? 1 : 0;
}
// Scene::SetTransformWorldMatrix / SetTransformLocalMatrix (engine/component.cpp:265-281, :236-248) on the transform of a node
void hgref_set_transform_matrix(void *h, int node, const float *m12, int is_local) {
	auto *c = static_cast<RefCtx *>(h);
	const auto ref = c->scene->GetNodeTransformRef(c->nodes[node]);
	if (is_local)
		c->scene->SetTransformLocalMatrix(ref, mat4_from(m12));
	else
		c->scene->SetTransformWorldMatrix(ref, mat4_from(m12));
}
#ifndef HG_FACADE // reference-internal host math (the facade computes these on the device only)
void hgref_minmax_transform(const float *m12, const float *mm6, float *out6) {
	const MinMax o = mat4_from(m12) * MinMax(Vec3(mm6[0], mm6[1], mm6[2]), Vec3(mm6[3], mm6[4], mm6[5]));
	memcpy(out6, &o.mn, 12);
	memcpy(out6 + 3, &o.mx, 12);
}
// restated: ComputeModelDisplayListSortKeys, engine/scene_forward_pipeline.cpp:438-483 (that TU drags the whole forward pipeline in, so its
// loop is repeated here with the reference's own operator*(Mat4,Mat4), operator*(Mat4,Vec3) and ComputeSortKey, render_pipeline.cpp:89-94)
// over the list left by the last hgref_cull. view12 = ViewState::view.
int hgref_sort_keys(void *h, const float *view12, uint32_t *out) {
	auto *c = static_cast<RefCtx *>(h);
	Mat4 view;
	memcpy(&view, view12, 48);
	const auto &mtxs = c->scene->GetTransformWorldMatrices();
	size_t n = 0;
	for (const auto &dl : c->culled) {
		const Mat4 mdl_view_mtx = view * mtxs[dl.mtx_idx];
		const auto &bound = c->resources.models.Get_unsafe_(dl.mdl_idx).bounds[dl.lst_idx];
		const Vec3 vtx[8] = {{bound.mn.x, bound.mn.y, bound.mn.z}, {bound.mx.x, bound.mn.y, bound.mn.z}, {bound.mx.x, bound.mx.y, bound.mn.z},
			{bound.mn.x, bound.mx.y, bound.mn.z}, {bound.mn.x, bound.mn.y, bound.mx.z}, {bound.mx.x, bound.mn.y, bound.mx.z},
			{bound.mx.x, bound.mx.y, bound.mx.z}, {bound.mn.x, bound.mx.y, bound.mx.z}};
		float closest = std::numeric_limits<float>::max();
		for (auto &v : vtx) {
			const Vec3 in_view = mdl_view_mtx * v;
			if (in_view.z < closest)
				closest = in_view.z;
		}
		if (closest < 0.f)
			closest = 0.f;
		if (out)
			out[n] = ComputeSortKey(closest);
		++n;
	}
	return int(n);
}
int hgref_test_visibility(const float *frustum24, const float *mm6) {
	Frustum f;
	memcpy(f.data(), frustum24, 96);
	return int(TestVisibility(f, MinMax(Vec3(mm6[0], mm6[1], mm6[2]), Vec3(mm6[3], mm6[4], mm6[5]))));
}
void hgref_quat_to_euler(const float *q4, float *out3) { // scene.cpp:2172 : ToEuler(Normalize(rot))
	const Vec3 e = ToEuler(Normalize(Quaternion(q4[0], q4[1], q4[2], q4[3])));
	memcpy(out3, &e, 12);
}
// Evaluate(AnimTrackHermiteT<Vec3>) animation.h:155-169; returns kf interval through GetIntervalKeys :109-128
int hgref_eval_vec3_track(int n_keys, const int64_t *t, const float *v3, const float *tension, const float *bias, int64_t at, float *out3, int *kf) {
	AnimTrackHermiteT<Vec3> trk;
	for (int i = 0; i < n_keys; ++i) {
		AnimKeyHermiteT<Vec3> k;
		k.t = t[i];
		k.v = Vec3(v3[i * 3], v3[i * 3 + 1], v3[i * 3 + 2]);
		k.tension = tension ? tension[i] : 0.f;
		k.bias = bias ? bias[i] : 0.f;
		trk.keys.push_back(k);
	}
	Vec3 v(out3[0], out3[1], out3[2]);
	const bool r = Evaluate(trk, at, v);
	memcpy(out3, &v, 12);
	if (kf) {
		kf[0] = kf[1] = -1;
		kf[2] = GetIntervalKeys<AnimTrackHermiteT<Vec3>, Vec3>(trk, at, kf[0], kf[1]) ? 1 : 0;
	}
	return r ? 1 : 0;
}
int hgref_eval_quat_track(int n_keys, const int64_t *t, const float *v4, int64_t at, float *out4) {
	AnimTrackT<Quaternion> trk;
	for (int i = 0; i < n_keys; ++i)
		trk.keys.push_back({t[i], Quaternion(v4[i * 4], v4[i * 4 + 1], v4[i * 4 + 2], v4[i * 4 + 3])});
	Quaternion q(out4[0], out4[1], out4[2], out4[3]);
	const bool r = Evaluate(trk, at, q);
	memcpy(out4, &q, 16);
	return r ? 1 : 0;
}
#endif
float hgref_sinf(float x) { return sinf(x); } // the libm the reference links (matrix3.cpp:199-204)
float hgref_cosf(float x) { return cosf(x); }

// ---- CPU baseline timing (bench.py cpu_baseline / --impl reference). One frame = the reference's own
// Scene::Update + GetModelDisplayLists + copy + CullModelDisplayLists(opaque) [+ transparent], single thread (the path has no threads).
// out4 = seconds spent in {Update, GetModelDisplayLists, Cull(incl. copy), bones}, summed over frames. Returns visible count of last frame.
int hgref_time_frames(void *h, int frames, int64_t dt, const float *frustum24, int with_bones, double *out4) {
	auto *c = static_cast<RefCtx *>(h);
	Frustum f;
	memcpy(f.data(), frustum24, 96);
	out4[0] = out4[1] = out4[2] = out4[3] = 0;
	int visible = 0;
	std::vector<float> bones;
	for (int i = 0; i < frames; ++i) {
		const double t0 = now_s();
		c->scene->Update(dt);
		const double t1 = now_s();
		c->scene->GetModelDisplayLists(c->opaque, c->transparent, c->opaque_skinned, c->transparent_skinned, c->resources);
		const double t2 = now_s();
		c->culled = c->opaque;
		CullModelDisplayLists(f, c->culled, c->scene->GetTransformWorldMatrices(), c->resources);
		visible = int(c->culled.size());
		if (!c->transparent.empty()) {
			auto tr = c->transparent;
			CullModelDisplayLists(f, tr, c->scene->GetTransformWorldMatrices(), c->resources);
			visible += int(tr.size());
		}
		const double t3 = now_s();
		if (with_bones) {
			const int n = hgref_bone_matrices(h, 2, 0, nullptr);
			bones.resize(size_t(n) * 16);
			hgref_bone_matrices(h, 2, 0, bones.data());
		}
		const double t4 = now_s();
		out4[0] += t1 - t0;
		out4[1] += t2 - t1;
		out4[2] += t3 - t2;
		out4[3] += t4 - t3;
	}
	return visible;
}

#ifdef HG_FACADE
// ---- the product's own per-frame API, timed end to end (bench.py `e2e_facade`): SetCullingFrustums once, then every frame
// Scene::Update(dt) (visibility evaluated inside the update pass) + GetVisibleModelDisplayLists = the std::vector<ModelDisplayList>
// that GetModelDisplayLists + CullModelDisplayLists leave behind in the reference (render_pipeline.cpp:2098-2104), materialised on the host.
// out3 = seconds in {Update, GetVisibleModelDisplayLists (opaque + transparent), total wall}; returns the visible count of the last frame.
int hgfacade_time_fused_frames(void *h, int frames, int64_t dt, const float *frustum24, double *out3) {
	auto *c = static_cast<RefCtx *>(h);
	Frustum f;
	memcpy(f.data(), frustum24, 96);
	c->scene->SetCullingFrustums({f}, c->resources);
	out3[0] = out3[1] = out3[2] = 0;
	int visible = 0;
	std::vector<ModelDisplayList> tr;
	const double w0 = now_s();
	for (int i = 0; i < frames; ++i) {
		const double t0 = now_s();
		c->scene->Update(dt);
		const double t1 = now_s();
		c->scene->GetVisibleModelDisplayLists(0, false, c->culled, c->resources);
		c->scene->GetVisibleModelDisplayLists(0, true, tr, c->resources);
		const double t2 = now_s();
		visible = int(c->culled.size() + tr.size());
		out3[0] += t1 - t0;
		out3[1] += t2 - t1;
	}
	out3[2] = now_s() - w0;
	return visible;
}
// seconds spent in GetVisibleModelDisplayLists since the last call of this function: device round trip, vector sizing, copy
void hgfacade_visible_breakdown(double *out3) {
	for (int i = 0; i < 3; ++i)
		out3[i] = hg::g_visible_breakdown[i], hg::g_visible_breakdown[i] = 0;
}
// one process per GPU: scenes created after this call live on `device` (Scene::SetDefaultDevice)
void hgfacade_set_default_device(int device) { Scene::SetDefaultDevice(device); }
void hgfacade_set_default_append_slack(int slots) { Scene::SetDefaultAppendSlack(slots); }
// kernels the scene's device context has launched so far (tests: an in-place edit must not re-sort, rebuild or re-capture)
uint64_t hgfacade_launch_count(void *h) {
	hgb_ctx *ctx = static_cast<RefCtx *>(h)->scene->GetDeviceContext();
	return ctx ? hgb_launch_count(ctx) : 0;
}
uint64_t hgfacade_rebuild_count(void *h) {
	hgb_ctx *ctx = static_cast<RefCtx *>(h)->scene->GetDeviceContext();
	return ctx ? hgb_rebuild_count(ctx) : 0;
}
// the list of the last fused frame, as mtx_idx per entry (parity check of the leg above)
int hgfacade_read_visible(void *h, uint32_t *out_mtx, int capacity) {
	auto *c = static_cast<RefCtx *>(h);
	const int n = int(c->culled.size());
	for (int i = 0; i < n && i < capacity; ++i)
		out_mtx[i] = c->culled[i].mtx_idx;
	return n;
}
#endif

} // extern "C"
