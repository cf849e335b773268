// hg::Scene facade, second translation unit: the host-side components and graph queries around the scene-update path that carry no
// device work of their own — node names and path lookup, Light components (animation targets), scene instances (NF_InstanceDisabled
// propagation, instance views, on-instantiate anims) and the free Create* helpers of engine/scene.h:915-933. Same member names,
// argument meaning and warn-and-neutral-value error behaviour as the reference (file:line given per function).
#include "engine/scene.h"
#include "foundation/log.h"
#include "hg_scene_impl.h"
#include "hgb200.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>

namespace hg {

const SceneAnimRef InvalidSceneAnimRef;

// ---- handles -----------------------------------------------------------------------------------------------------------------------
std::string Node::GetName() const { return scene ? scene->GetNodeName(ref) : std::string(); }
void Node::SetName(const std::string &v) { if (scene) scene->SetNodeName(ref, v); }
bool Node::IsItselfEnabled() const { return scene && scene->IsNodeItselfEnabled(ref); }
Light Node::GetLight() const { return scene ? scene->GetNodeLight(ref) : Light{}; }
void Node::SetLight(const Light &l) { if (scene) scene->SetNodeLight(ref, l.ref); }
Instance Node::GetInstance() const { return scene ? scene->GetNodeInstance(ref) : Instance{}; }
void Node::SetInstance(const Instance &i) { if (scene) scene->SetNodeInstance(ref, i.ref); }
bool Node::IsInstantiatedBy() const { return scene && scene->IsInstantiatedBy(ref) != InvalidNodeRef; }

bool Light::IsValid() const { return scene && scene->IsValidLightRef(ref); }
LightType Light::GetType() const { return scene ? scene->GetLightType(ref) : LT_Point; }
void Light::SetType(LightType v) { if (scene) scene->SetLightType(ref, v); }
Color Light::GetDiffuseColor() const { return scene ? scene->GetLightDiffuseColor(ref) : Color(); }
void Light::SetDiffuseColor(const Color &v) { if (scene) scene->SetLightDiffuseColor(ref, v); }
float Light::GetDiffuseIntensity() const { return scene ? scene->GetLightDiffuseIntensity(ref) : 0.f; }
void Light::SetDiffuseIntensity(float v) { if (scene) scene->SetLightDiffuseIntensity(ref, v); }
Color Light::GetSpecularColor() const { return scene ? scene->GetLightSpecularColor(ref) : Color(); }
void Light::SetSpecularColor(const Color &v) { if (scene) scene->SetLightSpecularColor(ref, v); }
float Light::GetSpecularIntensity() const { return scene ? scene->GetLightSpecularIntensity(ref) : 0.f; }
void Light::SetSpecularIntensity(float v) { if (scene) scene->SetLightSpecularIntensity(ref, v); }
float Light::GetRadius() const { return scene ? scene->GetLightRadius(ref) : 0.f; }
void Light::SetRadius(float v) { if (scene) scene->SetLightRadius(ref, v); }

bool Instance::IsValid() const { return scene && scene->IsValidInstanceRef(ref); }
std::string Instance::GetPath() const { return scene ? scene->GetInstancePath(ref) : std::string(); }
void Instance::SetPath(const std::string &path) { if (scene) scene->SetInstancePath(ref, path); }
void Instance::SetOnInstantiateAnim(const std::string &anim) { if (scene) scene->SetOnInstantiateAnim(ref, anim); }
void Instance::SetOnInstantiateAnimLoopMode(AnimLoopMode loop_mode) { if (scene) scene->SetOnInstantiateAnimLoopMode(ref, loop_mode); }
std::string Instance::GetOnInstantiateAnim() const { return scene ? scene->GetOnInstantiateAnim(ref) : std::string(); }

// ---- node names and path lookup (scene.cpp:522-579) ---------------------------------------------------------------------------------
std::string Scene::GetNodeName(NodeRef ref) const {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node reference");
		return {};
	}
	return nodes[ref.idx].name;
}
void Scene::SetNodeName(NodeRef ref, const std::string &v) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node reference");
		return;
	}
	nodes[ref.idx].name = v;
}
Node Scene::GetNode(const std::string &name) const {
	for (uint32_t i = nodes.first(); i != nodes.invalid_idx; i = nodes.next(i))
		if (nodes[i].name == name)
			return {const_cast<Scene *>(this), nodes.get_ref(i)};
	return {const_cast<Scene *>(this), InvalidNodeRef};
}
bool Scene::IsRoot(NodeRef ref) const { // scene.cpp: no transform, or a transform without a valid parent
	if (!nodes.is_valid(ref))
		return false;
	const ComponentRef t = nodes[ref.idx].transform;
	return !transforms.is_valid(t) || !nodes.is_valid(transforms[t.idx].parent);
}
std::vector<NodeRef> Scene::GetNodeChildRefs(NodeRef ref) const {
	std::vector<NodeRef> out;
	for (uint32_t i = nodes.first(); i != nodes.invalid_idx; i = nodes.next(i)) {
		const ComponentRef t = nodes[i].transform;
		if (transforms.is_valid(t) && transforms[t.idx].parent == ref)
			out.push_back(nodes.get_ref(i));
	}
	return out;
}
NodeRef Scene::GetNodeEx_(const std::vector<NodeRef> &refs, const std::string &path) const {
	int mode = 0; // 0: last element, 1: ':' enters the instance of the named node, 2: '/' descends into its children
	size_t s = 0;
	for (; s < path.size(); ++s)
		if (path[s] == ':') {
			mode = 1;
			break;
		} else if (path[s] == '/') {
			mode = 2;
			break;
		}
	const std::string name = path.substr(0, s), remainder = s < path.size() ? path.substr(s + 1) : std::string();
	for (const NodeRef &ref : refs)
		if (nodes[ref.idx].name == name) {
			if (mode == 0)
				return ref; // look no further
			if (mode == 1) {
				const auto i = node_instance_view.find(ref);
				if (i == node_instance_view.end())
					return InvalidNodeRef; // not an instance
				std::vector<NodeRef> roots;
				for (const NodeRef &n : i->second.nodes)
					if (IsChildOf(n, ref))
						roots.push_back(n);
				return GetNodeEx_(roots, remainder);
			}
			return GetNodeEx_(GetNodeChildRefs(ref), remainder);
		}
	return InvalidNodeRef;
}
Node Scene::GetNodeEx(const std::string &path) const {
	std::vector<NodeRef> roots;
	for (uint32_t i = nodes.first(); i != nodes.invalid_idx; i = nodes.next(i))
		if (IsRoot(nodes.get_ref(i)))
			roots.push_back(nodes.get_ref(i));
	return GetNode(GetNodeEx_(roots, path));
}
std::vector<Node> Scene::GetNodes() const {
	std::vector<Node> out;
	for (uint32_t i = nodes.first(); i != nodes.invalid_idx; i = nodes.next(i))
		if (!(nodes[i].flags & NF_Instantiated))
			out.push_back({const_cast<Scene *>(this), nodes.get_ref(i)});
	return out;
}
std::vector<Node> Scene::GetAllNodes() const {
	std::vector<Node> out;
	for (uint32_t i = nodes.first(); i != nodes.invalid_idx; i = nodes.next(i))
		out.push_back({const_cast<Scene *>(this), nodes.get_ref(i)});
	return out;
}

// ---- lights (scene.h:363-423, component.cpp): plain host-side values ------------------------------------------------------------------
#define HG_LGT(ret)                                                                                                                                            \
	if (!lights.is_valid(ref)) {                                                                                                                              \
		warn("Invalid light");                                                                                                                                \
		return ret;                                                                                                                                           \
	}
Light Scene::CreateLight() { return {this, lights.add_ref(Light_{})}; }
void Scene::DestroyLight(ComponentRef ref) { lights.remove_ref(ref); }
LightType Scene::GetLightType(ComponentRef ref) const { HG_LGT(LT_Point) return lights[ref.idx].type; }
void Scene::SetLightType(ComponentRef ref, LightType v) { HG_LGT() lights[ref.idx].type = v; }
Color Scene::GetLightDiffuseColor(ComponentRef ref) const { HG_LGT(Color()) return lights[ref.idx].diffuse; }
void Scene::SetLightDiffuseColor(ComponentRef ref, const Color &v) { HG_LGT() lights[ref.idx].diffuse = v; }
float Scene::GetLightDiffuseIntensity(ComponentRef ref) const { HG_LGT(0.f) return lights[ref.idx].diffuse_intensity; }
void Scene::SetLightDiffuseIntensity(ComponentRef ref, float v) { HG_LGT() lights[ref.idx].diffuse_intensity = v; }
Color Scene::GetLightSpecularColor(ComponentRef ref) const { HG_LGT(Color()) return lights[ref.idx].specular; }
void Scene::SetLightSpecularColor(ComponentRef ref, const Color &v) { HG_LGT() lights[ref.idx].specular = v; }
float Scene::GetLightSpecularIntensity(ComponentRef ref) const { HG_LGT(0.f) return lights[ref.idx].specular_intensity; }
void Scene::SetLightSpecularIntensity(ComponentRef ref, float v) { HG_LGT() lights[ref.idx].specular_intensity = v; }
float Scene::GetLightRadius(ComponentRef ref) const { HG_LGT(0.f) return lights[ref.idx].radius; }
void Scene::SetLightRadius(ComponentRef ref, float v) { HG_LGT() lights[ref.idx].radius = v; }
#undef HG_LGT

Light Scene::CreatePointLight(float radius, const Color &diffuse, float diffuse_intensity, const Color &specular, float specular_intensity, float priority,
	LightShadowType shadow_type, float shadow_bias) {
	Light_ l;
	l.type = LT_Point;
	l.radius = radius;
	l.diffuse = diffuse, l.diffuse_intensity = diffuse_intensity;
	l.specular = specular, l.specular_intensity = specular_intensity;
	l.priority = priority, l.shadow_type = shadow_type, l.shadow_bias = shadow_bias;
	return {this, lights.add_ref(l)};
}
std::vector<Node> Scene::GetLights() const {
	std::vector<Node> out;
	for (uint32_t i = nodes.first(); i != nodes.invalid_idx; i = nodes.next(i))
		if (lights.is_valid(nodes[i].light))
			out.push_back({const_cast<Scene *>(this), nodes.get_ref(i)});
	return out;
}
ComponentRef Scene::GetNodeLightRef(NodeRef ref) const {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return InvalidComponentRef;
	}
	return nodes[ref.idx].light;
}
void Scene::SetNodeLight(NodeRef ref, ComponentRef cref) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return;
	}
	nodes[ref.idx].light = cref;
	impl->anims_dirty = true; // light targets of playing anims are resolved at bind time
}

// ---- enable / disable with propagation into instantiated scenes (scene.cpp:472-503) -------------------------------------------------
bool Scene::IsNodeItselfEnabled(NodeRef ref) const {
	if (!nodes.is_valid(ref))
		return false;
	RefreshFlagsFromDevice();
	return !(nodes[ref.idx].flags & NF_Disabled);
}
void Scene::EnableNode_(NodeRef ref, bool through_instance) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node reference");
		return;
	}
	nodes[ref.idx].flags &= through_instance ? ~NF_InstanceDisabled : ~NF_Disabled;
	impl->flag_nodes.push_back(ref.idx);
	if (nodes[ref.idx].flags & (NF_Disabled | NF_InstanceDisabled)) // instance content comes back only when the host is fully enabled
		return;
	const auto i = node_instance_view.find(ref);
	if (i != node_instance_view.end())
		for (const NodeRef &n : i->second.nodes)
			EnableNode_(n, true);
}
void Scene::DisableNode_(NodeRef ref, bool through_instance) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node reference");
		return;
	}
	nodes[ref.idx].flags |= through_instance ? NF_InstanceDisabled : NF_Disabled;
	impl->flag_nodes.push_back(ref.idx);
	const auto i = node_instance_view.find(ref);
	if (i != node_instance_view.end())
		for (const NodeRef &n : i->second.nodes)
			DisableNode_(n, true);
}
void Scene::EnableNode(NodeRef ref) {
	RefreshFlagsFromDevice(); // device-evaluated "Enable" tracks first: this explicit call lands after them
	EnableNode_(ref, false); // the entries of the flipped nodes change their flag in place at the next sync (flag_nodes)
	impl->enable_state_push = true;
}
void Scene::DisableNode(NodeRef ref) {
	RefreshFlagsFromDevice();
	DisableNode_(ref, false);
	impl->enable_state_push = true;
}

// ---- scene instances (scene.h:437-476, scene.cpp:1190-1430) --------------------------------------------------------------------------
#define HG_INST(ret)                                                                                                                                           \
	if (!instances.is_valid(ref)) {                                                                                                                           \
		warn("Invalid instance");                                                                                                                             \
		return ret;                                                                                                                                           \
	}
Instance Scene::CreateInstance() { return {this, instances.add_ref(Instance_{})}; }
Instance Scene::CreateInstance(const std::string &path) {
	Instance i = CreateInstance();
	i.SetPath(path);
	return i;
}
void Scene::DestroyInstance(ComponentRef ref) { instances.remove_ref(ref); }
void Scene::SetInstancePath(ComponentRef ref, const std::string &path) { HG_INST() instances[ref.idx].name = path; }
std::string Scene::GetInstancePath(ComponentRef ref) const { HG_INST({}) return instances[ref.idx].name; }
void Scene::SetOnInstantiateAnim(ComponentRef ref, const std::string &anim) { HG_INST() instances[ref.idx].anim = anim; }
void Scene::SetOnInstantiateAnimLoopMode(ComponentRef ref, AnimLoopMode loop_mode) { HG_INST() instances[ref.idx].loop_mode = loop_mode; }
std::string Scene::GetOnInstantiateAnim(ComponentRef ref) const { HG_INST({}) return instances[ref.idx].anim; }
AnimLoopMode Scene::GetOnInstantiateAnimLoopMode(ComponentRef ref) const { HG_INST(ALM_Once) return instances[ref.idx].loop_mode; }
ScenePlayAnimRef Scene::GetOnInstantiatePlayAnimRef(ComponentRef ref) const { HG_INST(InvalidScenePlayAnimRef) return instances[ref.idx].play_anim_ref; }
#undef HG_INST

ComponentRef Scene::GetNodeInstanceRef(NodeRef ref) const {
	const auto i = node_instance.find(ref);
	return i != node_instance.end() ? i->second : InvalidComponentRef;
}
void Scene::SetNodeInstance(NodeRef ref, ComponentRef cref) {
	if (cref == InvalidComponentRef)
		node_instance.erase(ref);
	else
		node_instance[ref] = cref;
}
Instance Scene::GetNodeInstance(NodeRef ref) const {
	const ComponentRef c = GetNodeInstanceRef(ref);
	return c != InvalidComponentRef ? Instance{const_cast<Scene *>(this), c} : Instance{};
}
NodeRef Scene::IsInstantiatedBy(NodeRef ref) const {
	for (const auto &i : node_instance_view)
		for (const NodeRef &r : i.second.nodes)
			if (r == ref)
				return i.first;
	return InvalidNodeRef;
}
const SceneView &Scene::GetNodeInstanceSceneView(NodeRef ref) const {
	static const SceneView dummy;
	const auto i = node_instance_view.find(ref);
	if (i == node_instance_view.end()) {
		warn("No instance scene view on node");
		return dummy;
	}
	return i->second;
}
SceneAnimRef SceneView::GetSceneAnim(const Scene &scene, const std::string &name) const {
	for (const SceneAnimRef &r : scene_anims) {
		const SceneAnim *a = scene.GetSceneAnim(r);
		if (a && a->name == name)
			return r;
	}
	return InvalidSceneAnimRef;
}
SceneAnimRef Scene::GetSceneAnim(const std::string &name) const {
	for (uint32_t i = scene_anims.first(); i != scene_anims.invalid_idx; i = scene_anims.next(i))
		if (scene_anims[i].name == name)
			return scene_anims.get_ref(i);
	return InvalidSceneAnimRef;
}

// scene.cpp:1304-1348: the instance's scene file is loaded INTO this scene; what it brought is recorded as the node's scene view, flagged
// NF_Instantiated (and NF_InstanceDisabled when the host node is disabled) and parented to the host node
bool Scene::NodeSetupInstance_(NodeRef ref, const void *data, size_t size, const char *name, PipelineResources &resources, const PipelineInfo &pipeline,
	uint32_t flags, int recursion_level) {
	if (recursion_level > 4)
		return true;
	const auto i = node_instance.find(ref);
	if (i == node_instance.end())
		return false;
	const bool host_is_enabled = IsNodeEnabled(ref);
	if (!instances.is_valid(i->second))
		return true;
	LoadSceneContext ctx;
	ctx.recursion_level = recursion_level;
	if (!Load_binary(data, size, name, resources, pipeline, ctx, flags))
		return false;
	for (const NodeRef &n : ctx.view.nodes) {
		Node_ &node = nodes[n.idx];
		node.flags |= NF_Instantiated;
		if (!host_is_enabled)
			node.flags |= NF_InstanceDisabled;
		if (transforms.is_valid(node.transform) && transforms[node.transform.idx].parent == InvalidNodeRef)
			transforms[node.transform.idx].parent = ref;
	}
	for (const AnimRef &a : ctx.view.anims)
		anims[a.idx].flags |= AF_Instantiated;
	for (const SceneAnimRef &a : ctx.view.scene_anims)
		scene_anims[a.idx].flags |= SAF_Instantiated;
	node_instance_view[ref] = std::move(ctx.view);
	MarkBulkEdit();
	return true;
}
bool Scene::NodeSetupInstanceFromMemory(NodeRef ref, const void *data, size_t size, PipelineResources &resources, const PipelineInfo &pipeline, uint32_t flags,
	int recursion_level) {
	const ComponentRef c = GetNodeInstanceRef(ref);
	const std::string name = instances.is_valid(c) ? instances[c.idx].name : std::string();
	return NodeSetupInstance_(ref, data, size, name.c_str(), resources, pipeline, flags, recursion_level);
}
bool Scene::NodeSetupInstanceFromFile(NodeRef ref, PipelineResources &resources, const PipelineInfo &pipeline, uint32_t flags, int recursion_level) {
	const ComponentRef c = GetNodeInstanceRef(ref);
	if (c == InvalidComponentRef)
		return false;
	if (!instances.is_valid(c))
		return true;
	const std::string path = instances[c.idx].name;
	std::ifstream f(path, std::ios::binary);
	if (!f) {
		warn(("Cannot open instance scene '" + path + "'").c_str());
		return false;
	}
	const std::vector<char> blob((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
	return NodeSetupInstance_(ref, blob.data(), blob.size(), path.c_str(), resources, pipeline, flags, recursion_level);
}
void Scene::NodeDestroyInstance(NodeRef ref) { // scene.cpp:1288-1299
	NodeStopOnInstantiateAnim(ref);
	const auto i = node_instance_view.find(ref);
	if (i == node_instance_view.end()) {
		warn("Invalid node instance view");
		return;
	}
	for (const NodeRef &n : i->second.nodes)
		DestroyNode(n);
	for (const AnimRef &a : i->second.anims)
		DestroyAnim(a);
	for (const SceneAnimRef &a : i->second.scene_anims)
		DestroySceneAnim(a);
	node_instance_view.erase(i);
	MarkBulkEdit();
}
ScenePlayAnimRef Scene::NodeStartOnInstantiateAnim(NodeRef ref) { // scene.cpp:1350-1373
	NodeStopOnInstantiateAnim(ref);
	const auto i = node_instance.find(ref);
	if (i == node_instance.end() || !instances.is_valid(i->second))
		return InvalidScenePlayAnimRef;
	Instance_ &inst = instances[i->second.idx];
	if (inst.anim.empty())
		return InvalidScenePlayAnimRef;
	const auto v = node_instance_view.find(ref);
	if (v == node_instance_view.end())
		return InvalidScenePlayAnimRef;
	return inst.play_anim_ref = PlayAnim(v->second.GetSceneAnim(*this, inst.anim), inst.loop_mode);
}
void Scene::NodeStopOnInstantiateAnim(NodeRef ref) { // scene.cpp:1375-1388
	const auto i = node_instance.find(ref);
	if (i == node_instance.end() || !instances.is_valid(i->second))
		return;
	Instance_ &inst = instances[i->second.idx];
	if (play_anims.is_valid(inst.play_anim_ref))
		StopAnim(inst.play_anim_ref);
	inst.play_anim_ref = InvalidScenePlayAnimRef;
}

// ---- free helpers (scene.h:915-933, scene.cpp): node + transform decomposed from a matrix + one component -----------------------------
// Decompose (foundation/matrix4.cpp:208-252) + ToEuler YXZ (matrix3.cpp:91-137) on the host, in the reference's operation order and
// through the same libm (double atan2 narrowed to float): an edit-time helper, not part of the per-frame path.
static TransformTRS decompose_trs(const Mat4 &m) {
	TransformTRS o;
	o.pos = Vec3(m.m[0][3], m.m[1][3], m.m[2][3]);
	const float vx[3] = {m.m[0][0], m.m[1][0], m.m[2][0]}, vy[3] = {m.m[0][1], m.m[1][1], m.m[2][1]}, vz[3] = {m.m[0][2], m.m[1][2], m.m[2][2]};
	float sx = sqrtf(vx[0] * vx[0] + vx[1] * vx[1] + vx[2] * vx[2]);
	const float sy = sqrtf(vy[0] * vy[0] + vy[1] * vy[1] + vy[2] * vy[2]);
	const float sz = sqrtf(vz[0] * vz[0] + vz[1] * vz[1] + vz[2] * vz[2]);
	const float lx = vy[1] * vz[2] - vy[2] * vz[1], ly = vy[2] * vz[0] - vy[0] * vz[2], lz = vy[0] * vz[1] - vy[1] * vz[0];
	if (lx * vx[0] + ly * vx[1] + lz * vx[2] < 0.f)
		sx = -sx;
	o.scl = Vec3(sx, sy, sz);
	float r[3][3];
	for (int k = 0; k < 3; ++k) {
		r[k][0] = sx ? vx[k] / sx : (k == 0 ? 1.f : 0.f);
		r[k][1] = sy ? vy[k] / sy : (k == 1 ? 1.f : 0.f);
		r[k][2] = sz ? vz[k] / sz : (k == 2 ? 1.f : 0.f);
	}
	const float s1 = -r[1][2];
	const float c1 = sqrtf(r[1][1] * r[1][1] + r[1][0] * r[1][0]);
	float a0, a2;
	const float a1 = float(atan2(double(s1), double(c1)));
	if (c1 > 1.1920929e-07f) {
		a0 = float(atan2(double(r[0][2]), double(r[2][2])));
		a2 = float(atan2(double(r[1][0]), double(r[1][1])));
	} else {
		a0 = 0.f;
		const float sign_s1 = s1 < 0.f ? -1.f : (s1 > 0.f ? 1.f : 0.f);
		a2 = float(double(-sign_s1) * atan2(double(-r[2][0]), double(r[0][0])));
	}
	o.rot = Vec3(a1, a0, a2);
	return o;
}
static Node node_with_transform(Scene &scene, const Mat4 &mtx) {
	Node node = scene.CreateNode();
	const TransformTRS trs = decompose_trs(mtx);
	node.SetTransform(scene.CreateTransform(trs.pos, trs.rot, trs.scl));
	return node;
}
Node CreateObject(Scene &scene, const Mat4 &mtx, const ModelRef &model, std::vector<Material> materials) {
	Node node = node_with_transform(scene, mtx);
	node.SetObject(scene.CreateObject(model, std::move(materials)));
	return node;
}
Node CreatePointLight(Scene &scene, const Mat4 &mtx, float radius, const Color &diffuse, float diffuse_intensity, const Color &specular, float specular_intensity,
	float priority, LightShadowType shadow_type, float shadow_bias) {
	Node node = node_with_transform(scene, mtx);
	node.SetLight(scene.CreatePointLight(radius, diffuse, diffuse_intensity, specular, specular_intensity, priority, shadow_type, shadow_bias));
	return node;
}
Node CreateCamera(Scene &scene, const Mat4 &mtx, float znear, float zfar, float fov) {
	Node node = node_with_transform(scene, mtx);
	node.SetCamera(scene.CreateCamera(znear, zfar, fov));
	return node;
}

} // namespace hg

namespace hg {
Material &Scene::GetObjectMaterial(ComponentRef ref, size_t slot_idx) { // component.cpp: warn + dummy on an invalid ref / slot
	static Material dummy;
	if (!objects.is_valid(ref) || slot_idx >= objects[ref.idx].materials.size()) {
		warn("Invalid object or material slot");
		dummy = Material{};
		return dummy;
	}
	return objects[ref.idx].materials[slot_idx];
}
} // namespace hg
