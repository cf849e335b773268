// hg::Scene facade — the four render-pipeline types the scene-update path touches (harfang/engine/render_pipeline.h:38,220-224,
// 419-424,577-603) and the cull entry point (render_pipeline.h:584). bgfx handles are 16-bit placeholders exactly as in the
// reference's headless test fixture (harfang/tests/engine/scene.cpp:116-122); nothing here talks to a graphics API.
#pragma once
#include "foundation/generational_vector_list.h"
#include "foundation/hg_math_types.h"
#include <map>
#include <string>
#include <vector>

namespace bgfx {
struct IndexBufferHandle { uint16_t idx; };
struct VertexBufferHandle { uint16_t idx; };
} // namespace bgfx
#ifndef BGFX_INVALID_HANDLE
#define BGFX_INVALID_HANDLE {uint16_t(0xffff)}
#endif

namespace hg {

static const int max_skinned_model_matrix_count = 32; // render_pipeline.h:38

enum BlendMode { BM_Additive, BM_Alpha, BM_Darken, BM_Lighten, BM_Multiply, BM_Opaque, BM_Screen, BM_LinearBurn, BM_AlphaRGB_AddAlpha, BM_Undefined };

struct MaterialValue { // render_pipeline.h: a named uniform of a material; `value` holds 1..16 floats
	int type{3};      // bgfx::UniformType::Vec4
	std::vector<float> value;
	uint16_t count{1};
};
struct Material { // what the scene-update path reads: the blend mode (GetModelDisplayLists, render_pipeline.cpp:394-417) and the named
                  // values "Material.<slot>.<name>" vec4 tracks animate (scene.cpp:2199-2215)
	BlendMode blend_mode{BM_Opaque};
	std::map<std::string, MaterialValue> values;
};
inline void SetMaterialValue(Material &m, const char *name, const Vec4 &v) { m.values[name] = MaterialValue{3, {v.x, v.y, v.z, v.w}, 1}; }
inline BlendMode GetMaterialBlendMode(const Material &m) { return m.blend_mode; }
inline void SetMaterialBlendMode(Material &m, BlendMode mode) { m.blend_mode = mode; }

struct DisplayList { // render_pipeline.h:220-224
	bgfx::IndexBufferHandle index_buffer = BGFX_INVALID_HANDLE;
	bgfx::VertexBufferHandle vertex_buffer = BGFX_INVALID_HANDLE;
	std::vector<uint16_t> bones_table;
};

struct Model { // render_pipeline.h:419-424
	std::vector<MinMax> bounds;     // one per list
	std::vector<DisplayList> lists;
	std::vector<uint16_t> mats;     // material slot per list
	std::vector<Mat4> bind_pose;    // one per bone
};

struct ModelRef {
	gen_ref ref;
};
inline bool operator==(const ModelRef &a, const ModelRef &b) { return a.ref == b.ref; }
static const ModelRef InvalidModelRef;

// name -> gen_ref cache (harfang/engine/resource_cache.h). Index 0xffff is never handed out (uint16 mdl_idx in display lists).
template <typename T, typename Ref> class ResourceCache {
public:
	Ref Add(const std::string &name, const T &res) {
		auto i = name_to_ref.find(name);
		if (i != name_to_ref.end()) {
			resources[i->second.idx] = res;
			return {i->second};
		}
		const gen_ref r = resources.add_ref(res);
		name_to_ref[name] = r;
		return {r};
	}
	Ref Has(const std::string &name) const {
		auto i = name_to_ref.find(name);
		return i == name_to_ref.end() ? Ref{} : Ref{i->second};
	}
	bool IsValidRef(Ref ref) const { return resources.is_valid(ref.ref); }
	uint16_t GetValidatedRefIndex(Ref ref) const { return resources.is_valid(ref.ref) ? uint16_t(ref.ref.idx) : uint16_t(0xffff); }
	const T &Get_unsafe_(uint16_t idx) const { return idx != 0xffff && resources.is_used(idx) ? resources[idx] : dflt; }
	const T &Get(Ref ref) const { return resources.is_valid(ref.ref) ? resources[ref.ref.idx] : dflt; }
	size_t size() const { return resources.size(); }
	uint32_t first() const { return resources.first(); }
	uint32_t next(uint32_t i) const { return resources.next(i); }

private:
	generational_vector_list<T> resources;
	std::map<std::string, gen_ref> name_to_ref;
	T dflt{};
};

struct PipelineResources {
	ResourceCache<Model, ModelRef> models;
};

struct ModelDisplayList { // 16 B, render_pipeline.h:577-582
	const Material *mat;
	uint32_t mtx_idx;
	uint16_t mdl_idx, lst_idx;
};

struct SkinnedModelDisplayList { // render_pipeline.h:596-603
	const Material *mat;
	uint32_t mtx_idxs[max_skinned_model_matrix_count];
	uint32_t bones_idxs[max_skinned_model_matrix_count];
	uint16_t bone_count;
	uint16_t mdl_idx, lst_idx;
};

// Stable in-place removal of the entries whose transformed bounds are V_Outside (render_pipeline.cpp:2098-2104). `mtxs` must be the
// vector returned by Scene::GetTransformWorldMatrices() of a live scene: it identifies the device context holding the matrices.
// OPT-IN (not in the reference, which never culls skinned lists: scene_forward_pipeline.cpp:560-562): stable in-place removal of the
// skinned lists whose bound — the union over the list's bones of Model::bounds[lst_idx] carried through the bone's skinning matrix — is
// V_Outside. Nothing calls it by default.
void CullSkinnedModelDisplayLists(const Frustum &frustum, std::vector<SkinnedModelDisplayList> &display_lists, const std::vector<Mat4> &mtxs,
	const PipelineResources &resources);
void CullModelDisplayLists(const Frustum &frustum, std::vector<ModelDisplayList> &display_lists, const std::vector<Mat4> &mtxs, const PipelineResources &resources);

// The bone-matrix loop of _DrawSkinnedModelDisplayLists (render_pipeline.cpp:2157-2188) as a callable: 16 floats per (list, bone) in
// list order, to_bgfx layout (:59-62). previous = the previous-frame twin (:2177-2181).
size_t ComputeSkinMatrices(const std::vector<SkinnedModelDisplayList> &display_lists, const std::vector<Mat4> &mtxs, const PipelineResources &resources,
	std::vector<float> &out16, bool previous = false);

} // namespace hg
