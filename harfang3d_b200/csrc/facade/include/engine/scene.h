// hg::Scene facade — the reference's scene API for the scene-update path (harfang/engine/scene.h:173-901, node.h:36-380), same member
// names, argument meaning and warn-and-neutral-value error behaviour. The scene graph (nodes, components, refs, anims) lives on the
// host like in the reference; the per-frame members forward to the CUDA path through the C ABI of include/hgb200.h:
//   StorePreviousWorldMatrices / ReadyWorldMatrices / UpdatePlayingAnims / ComputeWorldMatrices / FixupPreviousWorldMatrices / Update
//   GetTransformWorldMatrices / GetPreviousTransformWorldMatrices (lazy device -> host), SetNodeWorldMatrix (preset)
// There is no CPU implementation of those members here: without a CUDA device they warn() and leave the matrices untouched.
#pragma once
#include "engine/animation.h"
#include "engine/render_pipeline.h"
#include "foundation/generational_vector_list.h"
#include "foundation/hg_math_types.h"
#include <map>
#include <string>
#include <vector>

struct hgb_ctx;

namespace hg {

using NodeRef = gen_ref;
using ComponentRef = gen_ref;
using AnimRef = gen_ref;
using SceneAnimRef = gen_ref;
using ScenePlayAnimRef = gen_ref;
extern const NodeRef InvalidNodeRef;
extern const ComponentRef InvalidComponentRef;
extern const ScenePlayAnimRef InvalidScenePlayAnimRef;
extern time_ns UnspecifiedAnimTime;

struct TransformTRS { // engine/node.h:36-38
	Vec3 pos{0, 0, 0}, rot{0, 0, 0}, scl{1, 1, 1};
};

class Scene;

// value-type handles (engine/node.h:66-67,282-380): a scene pointer + a generational ref; stale handles fail IsValid()
struct Transform {
	Scene *scene{nullptr};
	ComponentRef ref;
	bool IsValid() const;
	Vec3 GetPos() const;
	void SetPos(const Vec3 &v);
	Vec3 GetRot() const;
	void SetRot(const Vec3 &v);
	Vec3 GetScale() const;
	void SetScale(const Vec3 &v);
	void SetParent(const NodeRef &parent);
};
struct Object {
	Scene *scene{nullptr};
	ComponentRef ref;
	bool IsValid() const;
};
struct CameraZRange { // node.h:71-73
	float znear{0.01f}, zfar{1000.f};
};
struct Camera { // node.h:77-100: the view the culling frusta of a frame come from (Scene::ComputeCurrentCameraViewState)
	Scene *scene{nullptr};
	ComponentRef ref;
	bool IsValid() const;
	float GetZNear() const;
	void SetZNear(float v);
	float GetZFar() const;
	void SetZFar(float v);
	CameraZRange GetZRange() const;
	void SetZRange(CameraZRange v);
	float GetFov() const;
	void SetFov(float v);
	bool GetIsOrthographic() const;
	void SetIsOrthographic(bool v);
	float GetSize() const;
	void SetSize(float v);
};
enum LightType { LT_Point, LT_Spot, LT_Linear }; // node.h:135-136
enum LightShadowType { LST_None, LST_Map };
static const float default_shadow_bias = 0.0001f;
struct Light { // node.h:138-180: host-side component; colours and intensities are animation targets (scene.cpp:2183-2192)
	Scene *scene{nullptr};
	ComponentRef ref;
	bool IsValid() const;
	LightType GetType() const;
	void SetType(LightType v);
	Color GetDiffuseColor() const;
	void SetDiffuseColor(const Color &v);
	float GetDiffuseIntensity() const;
	void SetDiffuseIntensity(float v);
	Color GetSpecularColor() const;
	void SetSpecularColor(const Color &v);
	float GetSpecularIntensity() const;
	void SetSpecularIntensity(float v);
	float GetRadius() const;
	void SetRadius(float v);
};
struct Instance { // node.h: a node that instantiates another scene file under itself
	Scene *scene{nullptr};
	ComponentRef ref;
	bool IsValid() const;
	std::string GetPath() const;
	void SetPath(const std::string &path);
	void SetOnInstantiateAnim(const std::string &anim);
	void SetOnInstantiateAnimLoopMode(AnimLoopMode loop_mode);
	std::string GetOnInstantiateAnim() const;
};
struct Node {
	Scene *scene{nullptr};
	NodeRef ref;
	bool IsValid() const;
	uint32_t GetUid() const { return ref.idx; } // node.h:289
	std::string GetName() const;
	void SetName(const std::string &v);
	void Enable();
	void Disable();
	bool IsEnabled() const;
	bool IsItselfEnabled() const;
	Light GetLight() const;
	void SetLight(const Light &l);
	Instance GetInstance() const;
	void SetInstance(const Instance &i);
	bool IsInstantiatedBy() const;
	Transform GetTransform() const;
	void SetTransform(const Transform &t);
	Object GetObject() const;
	void SetObject(const Object &o);
	Camera GetCamera() const;
	void SetCamera(const Camera &c);
	Mat4 GetWorld() const;
	void SetWorld(const Mat4 &m);
	bool GetMinMax(const PipelineResources &resources, MinMax &minmax) const; // node.cpp:106-119 (object nodes; no instance views here)
};
inline bool operator==(const Node &a, const Node &b) { return a.scene == b.scene && a.ref == b.ref; }
// node.cpp:265-285: union of the nodes' world-space bounding boxes — a device reduction over the scene's current world matrices
bool GetNodesMinMax(const std::vector<Node> &nodes, const PipelineResources &resources, MinMax &minmax);

struct NodeAnim { // scene.h:104-107
	NodeRef node;
	AnimRef anim;
};
struct SceneAnim { // scene.h:109-118
	std::string name;
	time_ns t_start{0}, t_end{0};
	AnimRef scene_anim;
	std::vector<NodeAnim> node_anims;
	time_ns frame_duration{50000000}; // 20 fps
	uint8_t flags{0};
};

static const uint32_t NF_Disabled = 0x1, NF_InstanceDisabled = 0x2, NF_Instantiated = 0x4; // scene.h:166-170
static const uint8_t SAF_Instantiated = 0x10;

// engine/load_save_scene_flags.h:10-32
static const uint32_t LSSF_Nodes = 0x1, LSSF_Scene = 0x2, LSSF_Anims = 0x4, LSSF_KeyValues = 0x8, LSSF_Physics = 0x10, LSSF_Scripts = 0x20, LSSF_All = 0xffff;
static const uint32_t LSSF_DoNotLoadResources = 0x40000; // always in force here: resources are resolved by name, never read from disk

extern const SceneAnimRef InvalidSceneAnimRef;
struct SceneView { // scene.h:145-155
	std::vector<NodeRef> nodes;
	std::vector<AnimRef> anims;
	std::vector<SceneAnimRef> scene_anims;
	SceneAnimRef GetSceneAnim(const Scene &scene, const std::string &name) const; // scene.cpp: first scene anim of the view with that name
};
struct LoadSceneContext { // scene.h:157-161
	int recursion_level{};
	SceneView view;
	std::map<uint32_t, NodeRef> node_refs;
};
struct PipelineInfo {}; // render_pipeline.h: shader pipeline description — not used by the scene-update path
uint32_t GetSceneBinaryFormatVersion();

class Scene {
public:
	Scene();
	~Scene();
	Scene(const Scene &) = delete;
	Scene &operator=(const Scene &) = delete;

	size_t GarbageCollect(); // components no node refers to are removed (scene.cpp:92-237, transform / object / anim passes)

	// ---- nodes (scene.h:191-234) ----
	Node CreateNode(std::string name = {});
	void DestroyNode(NodeRef ref);
	void DestroyNode(const Node &node) { DestroyNode(node.ref); }
	bool IsValidNodeRef(NodeRef ref) const { return nodes.is_valid(ref); }
	Node GetNode(NodeRef ref) const { return nodes.is_valid(ref) ? Node{const_cast<Scene *>(this), ref} : Node{}; } // scene.cpp:421
	std::vector<Node> GetNodesWithObject() const; // GetNodesWithComponent(NCI_Object), scene.cpp:604-640
	void ReserveNodes(size_t count) { nodes.reserve(count); }
	void EnableNode(NodeRef ref);
	void DisableNode(NodeRef ref);
	bool IsNodeEnabled(NodeRef ref) const; // scene.h:227 (reads back device-evaluated "Enable" tracks first)
	size_t GetNodeCount() const { return nodes.size(); }
	bool IsChildOf(NodeRef node, NodeRef parent) const; // component.cpp:21-33
	bool IsNodeItselfEnabled(NodeRef ref) const; // NF_Disabled alone (scene.h:228)
	std::string GetNodeName(NodeRef ref) const;
	void SetNodeName(NodeRef ref, const std::string &v);
	Node GetNode(const std::string &name) const;                // scene.cpp:522-528: first node with that name
	Node GetNodeEx(const std::string &path) const;              // scene.cpp:530-579: "a/b" walks children, "a:b" enters an instance
	std::vector<NodeRef> GetNodeChildRefs(NodeRef ref) const;   // nodes whose transform's parent is `ref`
	bool IsRoot(NodeRef ref) const;                             // no transform, or no valid parent
	std::vector<Node> GetNodes() const;                         // instantiated nodes are left out (scene.cpp:583-595)
	std::vector<Node> GetAllNodes() const;

	// ---- lights (scene.h:363-423): host-side component, colours / intensities are animation targets ----
	Light CreateLight();
	void DestroyLight(ComponentRef ref);
	void DestroyLight(const Light &l) { DestroyLight(l.ref); }
	bool IsValidLightRef(ComponentRef ref) const { return lights.is_valid(ref); }
	LightType GetLightType(ComponentRef ref) const;
	void SetLightType(ComponentRef ref, LightType v);
	Color GetLightDiffuseColor(ComponentRef ref) const;
	void SetLightDiffuseColor(ComponentRef ref, const Color &v);
	float GetLightDiffuseIntensity(ComponentRef ref) const;
	void SetLightDiffuseIntensity(ComponentRef ref, float v);
	Color GetLightSpecularColor(ComponentRef ref) const;
	void SetLightSpecularColor(ComponentRef ref, const Color &v);
	float GetLightSpecularIntensity(ComponentRef ref) const;
	void SetLightSpecularIntensity(ComponentRef ref, float v);
	float GetLightRadius(ComponentRef ref) const;
	void SetLightRadius(ComponentRef ref, float v);
	Light CreatePointLight(float radius, const Color &diffuse = {1, 1, 1}, float diffuse_intensity = 1, const Color &specular = {1, 1, 1},
		float specular_intensity = 1, float priority = 0, LightShadowType shadow_type = LST_None, float shadow_bias = default_shadow_bias);
	std::vector<Node> GetLights() const;
	ComponentRef GetNodeLightRef(NodeRef ref) const;
	void SetNodeLight(NodeRef ref, ComponentRef cref);
	Light GetNodeLight(NodeRef ref) const { return {const_cast<Scene *>(this), GetNodeLightRef(ref)}; }
	void SetNodeLight(NodeRef ref, const Light &v) { SetNodeLight(ref, v.ref); }

	// ---- scene instances (scene.h:437-476; scene.cpp:1190-1430): a node that loads another scene under itself. Disabling the host node
	// propagates NF_InstanceDisabled to the instantiated nodes (EnableNode_/DisableNode_, scene.cpp:472-503); an Anim's
	// instance_anim_track plays the instantiated scene's anims at a nested clock (scene.cpp:2217-2268) ----
	Instance CreateInstance();
	Instance CreateInstance(const std::string &path);
	void DestroyInstance(ComponentRef ref);
	bool IsValidInstanceRef(ComponentRef ref) const { return instances.is_valid(ref); }
	void SetInstancePath(ComponentRef ref, const std::string &path);
	std::string GetInstancePath(ComponentRef ref) const;
	void SetOnInstantiateAnim(ComponentRef ref, const std::string &anim);
	void SetOnInstantiateAnimLoopMode(ComponentRef ref, AnimLoopMode loop_mode);
	std::string GetOnInstantiateAnim(ComponentRef ref) const;
	AnimLoopMode GetOnInstantiateAnimLoopMode(ComponentRef ref) const;
	ScenePlayAnimRef GetOnInstantiatePlayAnimRef(ComponentRef ref) const;
	ComponentRef GetNodeInstanceRef(NodeRef ref) const;
	void SetNodeInstance(NodeRef ref, ComponentRef cref);
	Instance GetNodeInstance(NodeRef ref) const;
	bool NodeSetupInstanceFromFile(NodeRef ref, PipelineResources &resources, const PipelineInfo &pipeline, uint32_t flags = LSSF_All, int recursion_level = 0);
	bool NodeSetupInstanceFromMemory(NodeRef ref, const void *data, size_t size, PipelineResources &resources, const PipelineInfo &pipeline, uint32_t flags = LSSF_All,
		int recursion_level = 0);
	void NodeDestroyInstance(NodeRef ref);
	ScenePlayAnimRef NodeStartOnInstantiateAnim(NodeRef ref);
	void NodeStopOnInstantiateAnim(NodeRef ref);
	const SceneView &GetNodeInstanceSceneView(NodeRef ref) const;
	NodeRef IsInstantiatedBy(NodeRef ref) const;
	SceneAnimRef GetSceneAnim(const std::string &name) const;
	const SceneAnim *GetSceneAnim(SceneAnimRef ref) const { return scene_anims.is_valid(ref) ? &scene_anims[ref.idx] : nullptr; }

	// ---- transforms (scene.h:242-284) ----
	Transform CreateTransform();
	Transform CreateTransform(const Vec3 &pos, const Vec3 &rot = {0, 0, 0}, const Vec3 &scl = {1, 1, 1}, NodeRef parent = {});
	void DestroyTransform(ComponentRef ref);
	void ReserveTransforms(size_t count) { transforms.reserve(count); }
	bool IsValidTransformRef(ComponentRef ref) const { return transforms.is_valid(ref); }
	Vec3 GetTransformPos(ComponentRef ref) const;
	void SetTransformPos(ComponentRef ref, const Vec3 &v);
	Vec3 GetTransformRot(ComponentRef ref) const;
	void SetTransformRot(ComponentRef ref, const Vec3 &v);
	Vec3 GetTransformScale(ComponentRef ref) const;
	void SetTransformScale(ComponentRef ref, const Vec3 &v);
	TransformTRS GetTransformTRS(ComponentRef ref) const;
	void SetTransformTRS(ComponentRef ref, const TransformTRS &v);
	NodeRef GetTransformParent(ComponentRef ref) const;
	void SetTransformParent(ComponentRef ref, const NodeRef &v); // rejects cycles like component.cpp:184-193
	ComponentRef GetNodeTransformRef(NodeRef ref) const;
	void SetNodeTransform(NodeRef ref, ComponentRef cref);
	void SetNodeTransform(NodeRef ref, const Transform &v) { SetNodeTransform(ref, v.ref); }
	Mat4 GetNodeWorldMatrix(NodeRef ref) const;            // scene.cpp:750-766
	void SetNodeWorldMatrix(NodeRef ref, const Mat4 &world); // scene.cpp:768-786: wins over ComputeWorldMatrices for this frame
	// component.cpp:236-281: the world row is written at once and the transform's TRS is refreshed from Decompose(local). Consecutive calls
	// are queued and reach the device as ONE batched hgb_set_matrices (sequential semantics kept) at the next read or frame.
	void SetTransformLocalMatrix(ComponentRef ref, const Mat4 &local);
	void SetTransformWorldMatrix(ComponentRef ref, const Mat4 &world);
	Mat4 GetTransformWorldMatrix(uint32_t transform_idx) const;         // component.cpp:250-256
	Mat4 GetPreviousTransformWorldMatrix(uint32_t transform_idx) const; // component.cpp:258-264

	// ---- cameras (scene.h:295-324,594-599): host-side component; its view state is built from the DEVICE-computed world matrix of
	// the camera node with the reference's own projection / frustum arithmetic (scene.cpp:240-263, render_pipeline.cpp:67-81) ----
	Camera CreateCamera();
	Camera CreateCamera(float znear, float zfar, float fov = 0.78539816339f); // Deg(45)
	Camera CreateOrthographicCamera(float znear, float zfar, float size = 1.f);
	void DestroyCamera(ComponentRef ref);
	void DestroyCamera(const Camera &c) { DestroyCamera(c.ref); }
	bool IsValidCameraRef(ComponentRef ref) const { return cameras.is_valid(ref); }
	float GetCameraZNear(ComponentRef ref) const;
	void SetCameraZNear(ComponentRef ref, float v);
	float GetCameraZFar(ComponentRef ref) const;
	void SetCameraZFar(ComponentRef ref, float v);
	CameraZRange GetCameraZRange(ComponentRef ref) const;
	void SetCameraZRange(ComponentRef ref, CameraZRange v);
	float GetCameraFov(ComponentRef ref) const;
	void SetCameraFov(ComponentRef ref, float v);
	float GetCameraSize(ComponentRef ref) const;
	void SetCameraSize(ComponentRef ref, float v);
	bool GetCameraIsOrthographic(ComponentRef ref) const;
	void SetCameraIsOrthographic(ComponentRef ref, const bool &v);
	void SetCameraCenterOffset(ComponentRef ref, const Vec2 &v);
	Vec2 GetCameraCenterOffset(ComponentRef ref) const;
	ComponentRef GetNodeCameraRef(NodeRef ref) const;
	void SetNodeCamera(NodeRef ref, ComponentRef cref);
	Camera GetNodeCamera(NodeRef ref) const { return {const_cast<Scene *>(this), GetNodeCameraRef(ref)}; }
	void SetNodeCamera(NodeRef ref, const Camera &v) { SetNodeCamera(ref, v.ref); }
	Node GetCurrentCamera() const { return {const_cast<Scene *>(this), current_camera}; }
	void SetCurrentCamera(NodeRef ref) { current_camera = ref; }
	void SetCurrentCamera(const Node &n) { current_camera = n.ref; }
	ViewState ComputeCurrentCameraViewState(const Vec2 &aspect_ratio) const { return ComputeCameraViewState(current_camera, aspect_ratio); }
	ViewState ComputeCameraViewState(NodeRef ref, const Vec2 &aspect_ratio) const;

	// ---- objects (scene.h:327-358) ----
	Object CreateObject();
	Object CreateObject(const ModelRef &model, std::vector<Material> materials = {});
	void DestroyObject(ComponentRef ref);
	bool IsValidObjectRef(ComponentRef ref) const { return objects.is_valid(ref); }
	void SetObjectModel(ComponentRef ref, const ModelRef &model);
	ModelRef GetObjectModel(ComponentRef ref) const;
	void SetObjectMaterialCount(ComponentRef ref, size_t count);
	void SetObjectMaterial(ComponentRef ref, size_t slot, Material mat);
	Material &GetObjectMaterial(ComponentRef ref, size_t slot_idx); // scene.h:342; an invalid ref / slot yields a scratch material
	size_t GetObjectBoneCount(ComponentRef ref) const;
	void SetObjectBoneCount(ComponentRef ref, size_t count);
	bool SetObjectBone(ComponentRef ref, size_t idx, NodeRef bone_node); // component.cpp:679-691
	NodeRef GetObjectBone(ComponentRef ref, size_t idx) const;
	ComponentRef GetNodeObjectRef(NodeRef ref) const;
	bool GetObjectMinMax(ComponentRef ref, const PipelineResources &resources, MinMax &minmax) const; // component.cpp:778-797
	bool GetMinMax(const PipelineResources &resources, MinMax &minmax) const;                         // scene.cpp:1982
	void SetNodeObject(NodeRef ref, ComponentRef cref);
	void SetNodeObject(NodeRef ref, const Object &v) { SetNodeObject(ref, v.ref); }

	// ---- per-frame path (scene.h:287-292,601-614) ----
	void StorePreviousWorldMatrices();
	void ReadyWorldMatrices();
	void ComputeWorldMatrices();
	void FixupPreviousWorldMatrices();
	void Update(time_ns dt);
	const std::vector<Mat4> &GetTransformWorldMatrices() const;
	const std::vector<Mat4> &GetPreviousTransformWorldMatrices() const;
	void GetModelDisplayLists(std::vector<ModelDisplayList> &out_opaque, std::vector<ModelDisplayList> &out_transparent,
		std::vector<SkinnedModelDisplayList> &out_opaque_skinned, std::vector<SkinnedModelDisplayList> &out_transparent_skinned,
		const PipelineResources &resources) const;

	// ---- animation (scene.h:620-663) ----
	AnimRef AddAnim(Anim anim);
	void DestroyAnim(AnimRef ref);
	bool IsValidAnim(AnimRef ref) const { return anims.is_valid(ref); }
	Anim *GetAnim(AnimRef ref);
	const Anim *GetAnim(AnimRef ref) const;
	SceneAnimRef AddSceneAnim(SceneAnim anim);
	void DestroySceneAnim(SceneAnimRef ref);
	SceneAnim *GetSceneAnim(SceneAnimRef ref);
	ScenePlayAnimRef PlayAnim(SceneAnimRef ref, AnimLoopMode loop_mode = ALM_Once, Easing easing = E_Linear, time_ns t_start = UnspecifiedAnimTime,
		time_ns t_end = UnspecifiedAnimTime, bool paused = false, float t_scale = 1.f);
	bool IsPlaying(ScenePlayAnimRef ref) const;
	void StopAnim(ScenePlayAnimRef ref);
	void StopAllAnims();
	void UpdatePlayingAnims(time_ns dt);

	// ---- scene binary (scene.h:676-677; format version 9) ----
	// Load_binary of the reference takes a Reader/Handle pair; here the file image is parsed from memory straight into the containers
	// (bulk, no per-node API calls) and the hierarchy reaches the device as one SoA upload. facade/src/hg_scene_load_binary.cpp
	bool Load_binary(const void *data, size_t size, const char *name, PipelineResources &resources, const PipelineInfo &pipeline, LoadSceneContext &ctx,
		uint32_t load_flags = LSSF_All);

	// ---- extensions of the B200 path (not in the reference) ----
	// Up to 8 frusta evaluated inside Update (fused with the world-matrix pass). GetVisibleModelDisplayLists then returns what
	// GetModelDisplayLists + CullModelDisplayLists(view) would, without a second pass over the matrices.
	void SetCullingFrustums(const std::vector<Frustum> &frustums, const PipelineResources &resources); // resources must outlive the scene's use of it
	bool GetVisibleModelDisplayLists(uint32_t view, bool transparent, std::vector<ModelDisplayList> &out, const PipelineResources &resources);
	hgb_ctx *GetDeviceContext() const { return ctx; }
	int GetDeviceIndex() const { return device; }
	static void SetDefaultDevice(int device);
	// Extension: scenes created after this call reserve `slots` free transform slots / display entries on the device (hgb_options.append_slack)
	// so that nodes spawned at the end of a live scene, and Enable / DisableNode, are applied in place. 0 (default) = off.
	static void SetDefaultAppendSlack(int slots);

private:
	friend bool GetNodesMinMax(const std::vector<Node> &, const PipelineResources &, MinMax &);
	friend void CullModelDisplayLists(const Frustum &, std::vector<ModelDisplayList> &, const std::vector<Mat4> &, const PipelineResources &);
	friend void CullSkinnedModelDisplayLists(const Frustum &, std::vector<SkinnedModelDisplayList> &, const std::vector<Mat4> &, const PipelineResources &);
	friend size_t ComputeSkinMatrices(const std::vector<SkinnedModelDisplayList> &, const std::vector<Mat4> &, const PipelineResources &, std::vector<float> &, bool);
	struct Impl;

	struct Node_ { // scene.h:698-702
		std::string name;
		uint32_t flags{0};
		ComponentRef transform, object, camera, light;
	};
	struct Light_ { // scene.h:759-774
		LightType type{LT_Point};
		LightShadowType shadow_type{LST_None};
		Color diffuse{1.f, 1.f, 1.f, 1.f};
		float diffuse_intensity{1.f};
		Color specular{1.f, 1.f, 1.f, 1.f};
		float specular_intensity{1.f};
		float radius{0.f};
		float inner_angle{0.5235987756f}, outer_angle{0.7853981634f}; // Deg(30), Deg(45)
		Vec4 pssm_split{10.f, 50.f, 100.f, 200.f};
		float priority{0.f};
		float shadow_bias{default_shadow_bias};
	};
	struct Instance_ { // scene.h: path of the instantiated scene + the anim started on instantiation
		std::string name, anim;
		AnimLoopMode loop_mode{ALM_Once};
		ScenePlayAnimRef play_anim_ref;
	};
	struct Camera_ { // scene.h:739-745
		CameraZRange zrange;
		float fov{0.69813170079f}; // Deg(40)
		bool ortho{false};
		float size{1.f};
		Vec2 center_offset{0.f, 0.f};
	};
	struct Transform_ { // scene.h:734-737
		TransformTRS TRS;
		NodeRef parent;
	};
	struct Object_ { // scene.h:750-757
		ModelRef model;
		std::vector<Material> materials;
		std::vector<NodeRef> bones;
	};
	struct ScenePlayAnim_ { // scene.h:884-895
		SceneAnimRef scene_anim;
		time_ns t_start, t_end, t;
		int8_t t_scale;
		AnimLoopMode loop_mode;
		bool paused;
		uint32_t play_slot; // index into the device's play clock array
	};

	generational_vector_list<Node_> nodes;
	generational_vector_list<Transform_> transforms;
	generational_vector_list<Object_> objects;
	generational_vector_list<Camera_> cameras;
	generational_vector_list<Light_> lights;
	generational_vector_list<Instance_> instances;
	std::map<NodeRef, ComponentRef> node_instance;    // scene.h: instance component of a node
	std::map<NodeRef, SceneView> node_instance_view;  // what the instance brought into this scene
	NodeRef current_camera;
	generational_vector_list<Anim> anims;
	generational_vector_list<SceneAnim> scene_anims;
	generational_vector_list<ScenePlayAnim_> play_anims;

	mutable std::vector<Mat4> transform_worlds, previous_transform_worlds; // host mirrors, refreshed lazily
	mutable bool worlds_valid{false}, previous_valid{false};
	mutable bool trs_host_stale{false}; // the device sampled tracks into the TRS: refresh before the host reads or edits them

	hgb_ctx *ctx{nullptr};
	int device{0};
	Impl *impl; // dirty tracking + staging (facade/src/hg_scene.cpp)

	void EnableNode_(NodeRef ref, bool through_instance);
	void DisableNode_(NodeRef ref, bool through_instance);
	NodeRef GetNodeEx_(const std::vector<NodeRef> &refs, const std::string &path) const;
	bool NodeSetupInstance_(NodeRef ref, const void *data, size_t size, const char *name, PipelineResources &resources, const PipelineInfo &pipeline, uint32_t flags,
		int recursion_level);
	void EvaluateHostTracks(); // light / camera / material values and nested instance anims of the playing anims (host-side single values)
	bool EnsureContext() const;
	void RefreshTRSFromDevice() const;
	void NoteFreshTransform(uint32_t idx);
	void RefreshFlagsFromDevice() const;
	void MarkBulkEdit(); // containers were filled behind the setters' back: everything is re-derived at the next sync
	void QueueMatrixSet(uint32_t slot, const Mat4 &m, bool is_local);
	void FlushMatrixSets() const;
	bool SyncEdits(const PipelineResources *resources);
};

// scene.h:915-933: a node with a transform (TRS decomposed from `mtx` by the device-free Decompose) and one component
Node CreateObject(Scene &scene, const Mat4 &mtx, const ModelRef &model, std::vector<Material> materials = {});
Node CreatePointLight(Scene &scene, const Mat4 &mtx, float radius, const Color &diffuse = {1, 1, 1}, float diffuse_intensity = 1,
	const Color &specular = {1, 1, 1}, float specular_intensity = 1, float priority = 0.f, LightShadowType shadow_type = LST_None,
	float shadow_bias = default_shadow_bias);
Node CreateCamera(Scene &scene, const Mat4 &mtx, float znear, float zfar, float fov = 0.69813170079f);

// scene.h:962-974
bool LoadSceneBinaryFromFile(const char *path, Scene &scene, PipelineResources &resources, const PipelineInfo &pipeline, LoadSceneContext &ctx,
	uint32_t flags = LSSF_All, bool debug = false);
bool LoadSceneBinaryFromMemory(const void *data, size_t size, const char *name, Scene &scene, PipelineResources &resources, const PipelineInfo &pipeline,
	LoadSceneContext &ctx, uint32_t flags = LSSF_All);

} // namespace hg
