// hg::Scene facade — host side of the drop-in boundary. The scene graph is kept on the host with the
// reference's semantics (slot reuse order, generational refs, warn-and-neutral-value errors); edits are tracked as dirty state and
// flushed to the device context (include/hgb200.h) at the start of a frame; the per-frame members only enqueue GPU work.
// Nothing in this file computes a world matrix, samples a key or tests a frustum: without a CUDA device those members warn and return.
#include "engine/scene.h"
#include "foundation/log.h"
#include "hg_scene_impl.h"
#include "hgb200.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <unordered_map>

namespace hg {

const NodeRef InvalidNodeRef;
const ComponentRef InvalidComponentRef;
const ScenePlayAnimRef InvalidScenePlayAnimRef;
time_ns UnspecifiedAnimTime = INT64_MIN; // scene.cpp: "unspecified" sentinel of PlayAnim's t_start / t_end

static int g_default_device = 0;
void Scene::SetDefaultDevice(int device) { g_default_device = device; }
static int g_default_append_slack = 0;
void Scene::SetDefaultAppendSlack(int slots) { g_default_append_slack = slots > 0 ? slots : 0; }

// CullModelDisplayLists / ComputeSkinMatrices receive the world-matrix vector, not the scene (render_pipeline.h:584): the vector's
// address identifies the scene whose device context holds the matrices.
static std::mutex g_registry_mutex;
static std::map<const void *, Scene *> g_registry;

// A few persistent host threads for the one per-frame host loop that is worth it: turning the visible INDEX list the device returns
// into the std::vector<ModelDisplayList> the caller asked for (1.6 M x 16 bytes per frame on the 16M-node batch: the records arrive
// contiguous in pinned memory — the device did the gather — and are copied into the caller's vector in 1 MB chunks). The scene itself stays single-threaded like the reference's
// (SURVEY.md §8b): the workers only ever run inside GetVisibleModelDisplayLists and touch nothing but its input and output arrays.
namespace {
class GatherPool {
public:
	static GatherPool &get() {
		static GatherPool p;
		return p;
	}
	// fn(chunk) for chunk in [0, n_chunks), the caller takes part; returns when all chunks are done
	void run(uint32_t n_chunks, const std::function<void(uint32_t)> &fn) {
		if (n_chunks <= 1 || workers.empty()) {
			for (uint32_t c = 0; c < n_chunks; ++c)
				fn(c);
			return;
		}
		uint64_t gen;
		{
			std::lock_guard<std::mutex> lock(m);
			gen = ++generation;
			job = &fn;
			total = n_chunks;
			done.store(0);
			ticket.store(gen << 32); // published last: a worker that sees this generation sees job / total / done of it
		}
		cv.notify_all();
		work(gen);
		std::unique_lock<std::mutex> lock(m);
		cv_done.wait(lock, [&] { return done.load() == n_chunks; });
	}

private:
	GatherPool() {
		const unsigned hw = std::thread::hardware_concurrency();
		const unsigned n = hw > 1 ? std::min(hw, 16u) - 1 : 0;
		for (unsigned i = 0; i < n; ++i)
			workers.emplace_back([this] { loop(); });
	}
	~GatherPool() {
		{
			std::lock_guard<std::mutex> lock(m);
			quit = true;
		}
		cv.notify_all();
		for (auto &t : workers)
			t.join();
	}
	// chunks are claimed by compare-and-swap on (generation, next): a worker still holding an older generation can neither take nor
	// skip a chunk of the current job
	void work(uint64_t gen) {
		for (;;) {
			uint64_t t = ticket.load();
			if ((t >> 32) != gen)
				return;
			const uint32_t c = uint32_t(t);
			if (c >= total)
				return;
			if (!ticket.compare_exchange_weak(t, t + 1))
				continue;
			(*job)(c);
			if (done.fetch_add(1) + 1 == total) {
				std::lock_guard<std::mutex> lock(m);
				cv_done.notify_all();
			}
		}
	}
	void loop() {
		uint64_t seen = 0;
		for (;;) {
			{
				std::unique_lock<std::mutex> lock(m);
				cv.wait(lock, [&] { return quit || generation != seen; });
				if (quit)
					return;
				seen = generation;
			}
			work(seen);
		}
	}
	std::vector<std::thread> workers;
	std::mutex m;
	std::condition_variable cv, cv_done;
	const std::function<void(uint32_t)> *job{nullptr};
	std::atomic<uint64_t> ticket{0};
	std::atomic<uint32_t> done{0};
	uint32_t total{0};
	uint64_t generation{0};
	bool quit{false};
};
} // namespace

static void check(int rc) {
	if (rc)
		warn(hgb_last_error());
}

Scene::Scene() : impl(new Impl) {
	device = g_default_device;
	impl->append_slack = g_default_append_slack;
	std::lock_guard<std::mutex> lock(g_registry_mutex);
	g_registry[&transform_worlds] = this;
	g_registry[&previous_transform_worlds] = this;
}

Scene::~Scene() {
	{
		std::lock_guard<std::mutex> lock(g_registry_mutex);
		g_registry.erase(&transform_worlds);
		g_registry.erase(&previous_transform_worlds);
	}
	if (ctx)
		hgb_destroy(ctx);
	delete impl;
}

bool Scene::EnsureContext() const {
	if (ctx)
		return true;
	Scene *self = const_cast<Scene *>(this);
	hgb_options opt{};
	opt.use_graph = 1;
	opt.append_slack = impl->append_slack; // the default in force when the scene was created
	self->ctx = hgb_create(device, &opt);
	if (!ctx) {
		if (!impl->warned_no_device)
			warn(hgb_last_error()); // no CPU fallback: the per-frame members do nothing without a device
		impl->warned_no_device = true;
		return false;
	}
	return true;
}

// ---- handles ---------------------------------------------------------------------------------------------------------------
bool Node::IsValid() const { return scene && scene->IsValidNodeRef(ref); }
void Node::Enable() { if (scene) scene->EnableNode(ref); }
void Node::Disable() { if (scene) scene->DisableNode(ref); }
bool Node::IsEnabled() const { return scene && scene->IsNodeEnabled(ref); }
Transform Node::GetTransform() const { return scene ? Transform{scene, scene->GetNodeTransformRef(ref)} : Transform{}; }
void Node::SetTransform(const Transform &t) { if (scene) scene->SetNodeTransform(ref, t.ref); }
Camera Node::GetCamera() const { return scene ? Camera{scene, scene->GetNodeCameraRef(ref)} : Camera{}; }
void Node::SetCamera(const Camera &c) { if (scene) scene->SetNodeCamera(ref, c.ref); }
Object Node::GetObject() const { return scene ? Object{scene, scene->GetNodeObjectRef(ref)} : Object{}; }
void Node::SetObject(const Object &o) { if (scene) scene->SetNodeObject(ref, o.ref); }
Mat4 Node::GetWorld() const { return scene ? scene->GetNodeWorldMatrix(ref) : Mat4::Identity; }
void Node::SetWorld(const Mat4 &m) { if (scene) scene->SetNodeWorldMatrix(ref, m); }
bool Transform::IsValid() const { return scene && scene->IsValidTransformRef(ref); }
Vec3 Transform::GetPos() const { return scene ? scene->GetTransformPos(ref) : Vec3(); }
void Transform::SetPos(const Vec3 &v) { if (scene) scene->SetTransformPos(ref, v); }
Vec3 Transform::GetRot() const { return scene ? scene->GetTransformRot(ref) : Vec3(); }
void Transform::SetRot(const Vec3 &v) { if (scene) scene->SetTransformRot(ref, v); }
Vec3 Transform::GetScale() const { return scene ? scene->GetTransformScale(ref) : Vec3(1, 1, 1); }
void Transform::SetScale(const Vec3 &v) { if (scene) scene->SetTransformScale(ref, v); }
void Transform::SetParent(const NodeRef &parent) { if (scene) scene->SetTransformParent(ref, parent); }
bool Object::IsValid() const { return scene && scene->IsValidObjectRef(ref); }

// ---- nodes -------------------------------------------------------------------------------------------------------------------
Node Scene::CreateNode(std::string name) {
	Node_ n;
	n.name = std::move(name);
	const NodeRef ref = nodes.add_ref(std::move(n));
	Impl::stamp(impl->node_epoch, ref.idx, impl->epoch);
	impl->touch_entries(ref.idx); // a node behind the last one of the entry tables only appends to them
	return {this, ref};
}

void Scene::DestroyNode(NodeRef ref) {
	if (!nodes.is_valid(ref))
		return;
	nodes.remove_ref(ref);
	impl->topology_dirty = true, impl->topology_reparent_only = impl->topology_append_only = false; // children of its transform become roots (a stale parent ref is a root, scene.cpp:268-272)
	impl->touch_entries_all();
	impl->anims_dirty = true;
}

// "Enable" tracks are evaluated on the device (k_sample_enable; EvaluateBoundAnim scene.cpp:2158-2162): the host's NF_Disabled of the
// tracked nodes is refreshed from it before anything on the host reads or edits node flags.
void Scene::RefreshFlagsFromDevice() const {
	if (!impl->flags_stale)
		return;
	impl->flags_stale = false;
	const size_t n = impl->enable_nodes.size();
	if (!ctx || !n)
		return;
	std::vector<uint32_t> slots(n);
	std::vector<uint8_t> dis(n);
	uint32_t cnt = 0;
	if (hgb_get_enable_state(ctx, uint32_t(n), slots.data(), dis.data(), &cnt)) {
		warn(hgb_last_error());
		return;
	}
	auto &nd = const_cast<generational_vector_list<Node_> &>(nodes);
	for (const auto &sn : impl->enable_nodes) {
		const auto it = std::lower_bound(slots.begin(), slots.begin() + cnt, sn.first);
		if (it == slots.begin() + cnt || *it != sn.first || !nodes.is_used(sn.second))
			continue;
		if (dis[it - slots.begin()])
			nd[sn.second].flags |= NF_Disabled;
		else
			nd[sn.second].flags &= ~NF_Disabled;
	}
}

bool Scene::IsNodeEnabled(NodeRef ref) const {
	if (!nodes.is_valid(ref))
		return false;
	RefreshFlagsFromDevice();
	return !(nodes[ref.idx].flags & (NF_Disabled | NF_InstanceDisabled));
}

bool Scene::IsChildOf(NodeRef node, NodeRef parent) const { // walk up from `node`; bounded by the node count against corrupted graphs
	size_t guard = nodes.capacity() + 1;
	while (guard--) {
		if (!nodes.is_valid(node))
			return false;
		const ComponentRef t = nodes[node.idx].transform;
		if (!transforms.is_valid(t))
			return false;
		const NodeRef up = transforms[t.idx].parent;
		if (up == parent)
			return true;
		node = up;
	}
	return false;
}

// ---- transforms --------------------------------------------------------------------------------------------------------------
// a transform created since the last sync: the first hierarchy edit of a batch that may be appended in place (SyncEdits)
void Scene::NoteFreshTransform(uint32_t idx) {
	impl->fresh_slots.push_back(idx);
	Impl::stamp(impl->slot_epoch, idx, impl->epoch);
	if (!impl->topology_dirty)
		impl->topology_append_only = true;
	impl->topology_reparent_only = false;
	impl->topology_dirty = true;
}

Transform Scene::CreateTransform() {
	const ComponentRef ref = transforms.add_ref(Transform_{});
	NoteFreshTransform(ref.idx);
	return {this, ref};
}

Transform Scene::CreateTransform(const Vec3 &pos, const Vec3 &rot, const Vec3 &scl, NodeRef parent) {
	Transform_ t;
	t.TRS.pos = pos, t.TRS.rot = rot, t.TRS.scl = scl;
	t.parent = parent;
	const ComponentRef ref = transforms.add_ref(t);
	NoteFreshTransform(ref.idx);
	return {this, ref};
}

void Scene::DestroyTransform(ComponentRef ref) {
	if (transforms.is_valid(ref)) {
		transforms.remove_ref(ref);
		impl->topology_dirty = true, impl->topology_reparent_only = impl->topology_append_only = false;
		impl->touch_entries_all();
		impl->anims_dirty = true;
	}
}

void Scene::MarkBulkEdit() {
	RefreshTRSFromDevice(); // before every slot is declared fresh
	impl->fresh_slots.clear();
	for (uint32_t i = transforms.first(); i != transforms.invalid_idx; i = transforms.next(i))
		impl->fresh_slots.push_back(i);
	impl->topology_dirty = true, impl->topology_reparent_only = impl->topology_append_only = false;
	impl->trs_all_dirty = true;
	impl->touch_entries_all();
	impl->anims_dirty = true;
}

// The device samples bound tracks straight into its TRS; the host copy is refreshed on the first host access after such a frame.
void Scene::RefreshTRSFromDevice() const {
	FlushMatrixSets();
	if (!trs_host_stale)
		return;
	trs_host_stale = false;
	// with a pending re-layout the device still holds the OLD topology: its slots are read back (sampled values must survive the
	// re-layout), except the ones (re)created on the host since
	const uint32_t n = std::min(uint32_t(transforms.capacity()), impl->device_slots);
	if (!ctx || !n)
		return;
	impl->staging.resize(size_t(n) * 9);
	if (hgb_get_trs(ctx, 0, n, impl->staging.data())) {
		warn(hgb_last_error());
		return;
	}
	std::vector<uint8_t> fresh(n, 0);
	for (uint32_t sl : impl->fresh_slots)
		if (sl < n)
			fresh[sl] = 1;
	auto &trs = const_cast<generational_vector_list<Transform_> &>(transforms);
	for (uint32_t i = transforms.first(); i != transforms.invalid_idx && i < n; i = transforms.next(i)) {
		if (fresh[i])
			continue;
		const float *s = impl->staging.data() + size_t(i) * 9;
		trs[i].TRS.pos = Vec3(s[0], s[1], s[2]), trs[i].TRS.rot = Vec3(s[3], s[4], s[5]), trs[i].TRS.scl = Vec3(s[6], s[7], s[8]);
	}
}

#define HG_GET_TRS(FIELD, DFLT)                                                                                                                                \
	if (!transforms.is_valid(ref)) {                                                                                                                           \
		warn("Invalid transform component");                                                                                                                   \
		return DFLT;                                                                                                                                           \
	}                                                                                                                                                          \
	RefreshTRSFromDevice();                                                                                                                                    \
	return transforms[ref.idx].TRS.FIELD;
#define HG_SET_TRS(FIELD)                                                                                                                                      \
	if (!transforms.is_valid(ref)) {                                                                                                                           \
		warn("Invalid transform component");                                                                                                                   \
		return;                                                                                                                                                \
	}                                                                                                                                                          \
	RefreshTRSFromDevice();                                                                                                                                    \
	transforms[ref.idx].TRS.FIELD = v;                                                                                                                         \
	impl->dirty_trs.push_back(ref.idx);

Vec3 Scene::GetTransformPos(ComponentRef ref) const { HG_GET_TRS(pos, Vec3()) }
void Scene::SetTransformPos(ComponentRef ref, const Vec3 &v) { HG_SET_TRS(pos) }
Vec3 Scene::GetTransformRot(ComponentRef ref) const { HG_GET_TRS(rot, Vec3()) }
void Scene::SetTransformRot(ComponentRef ref, const Vec3 &v) { HG_SET_TRS(rot) }
Vec3 Scene::GetTransformScale(ComponentRef ref) const { HG_GET_TRS(scl, Vec3(1, 1, 1)) }
void Scene::SetTransformScale(ComponentRef ref, const Vec3 &v) { HG_SET_TRS(scl) }

TransformTRS Scene::GetTransformTRS(ComponentRef ref) const {
	if (!transforms.is_valid(ref)) {
		warn("Invalid transform component");
		return {};
	}
	RefreshTRSFromDevice();
	return transforms[ref.idx].TRS;
}

void Scene::SetTransformTRS(ComponentRef ref, const TransformTRS &v) {
	if (!transforms.is_valid(ref)) {
		warn("Invalid transform component");
		return;
	}
	RefreshTRSFromDevice();
	transforms[ref.idx].TRS = v;
	impl->dirty_trs.push_back(ref.idx);
}

NodeRef Scene::GetTransformParent(ComponentRef ref) const {
	if (!transforms.is_valid(ref)) {
		warn("Invalid transform component");
		return InvalidNodeRef;
	}
	return transforms[ref.idx].parent;
}

void Scene::SetTransformParent(ComponentRef ref, const NodeRef &v) {
	if (!transforms.is_valid(ref)) {
		warn("Invalid transform component");
		return;
	}
	// refuse a parent whose own ancestry passes through this transform (component.cpp:21-33,184-193)
	NodeRef walk = v;
	size_t guard = nodes.capacity() + 1;
	while (nodes.is_valid(walk) && guard--) {
		const ComponentRef t = nodes[walk.idx].transform;
		if (t == ref) {
			warn("Cyclical reference detected");
			return;
		}
		if (!transforms.is_valid(t))
			break;
		walk = transforms[t.idx].parent;
	}
	transforms[ref.idx].parent = v;
	if (impl->topology_dirty && impl->topology_append_only && impl->fresh_slot(ref.idx))
		return; // a fresh transform finds its place: still an append (its parent is read at the sync)
	impl->topology_append_only = false;
	if (!impl->topology_dirty) // first hierarchy edit since the last sync: a candidate for the incremental path (hgb_set_parents)
		impl->topology_reparent_only = true, impl->reparented.clear();
	if (impl->topology_reparent_only)
		impl->reparented.push_back(ref.idx);
	impl->topology_dirty = true;
}

ComponentRef Scene::GetNodeTransformRef(NodeRef ref) const {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return InvalidComponentRef;
	}
	return nodes[ref.idx].transform;
}

void Scene::SetNodeTransform(NodeRef ref, ComponentRef cref) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return;
	}
	nodes[ref.idx].transform = cref;
	if (impl->fresh_node(ref.idx) && impl->fresh_slot(cref.idx) && impl->topology_dirty && impl->topology_append_only) {
		impl->touch_entries(ref.idx); // a node created a moment ago takes a transform created a moment ago: nothing refers to either yet
		return;
	}
	impl->topology_dirty = true, impl->topology_reparent_only = impl->topology_append_only = false; // children reach their parent's transform through the parent NODE (scene.cpp:268)
	impl->touch_entries_all();
	impl->anims_dirty = true;
}

Mat4 Scene::GetNodeWorldMatrix(NodeRef ref) const {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return Mat4::Identity;
	}
	const ComponentRef t = nodes[ref.idx].transform;
	if (!transforms.is_valid(t)) {
		warn("Invalid node transform");
		return Mat4::Identity;
	}
	const std::vector<Mat4> &w = GetTransformWorldMatrices();
	if (t.idx >= w.size()) {
		warn("Invalid node transform index");
		return Mat4::Identity;
	}
	return w[t.idx];
}

void Scene::SetNodeWorldMatrix(NodeRef ref, const Mat4 &world) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return;
	}
	const ComponentRef t = nodes[ref.idx].transform;
	if (!transforms.is_valid(t)) {
		warn("Invalid node transform");
		return;
	}
	// recorded as a preset of the frame being prepared: it wins over ComputeWorldMatrices and is the parent of the node's children
	impl->preset_slots.push_back(t.idx);
	impl->preset_mats.insert(impl->preset_mats.end(), &world.m[0][0], &world.m[0][0] + 12);
	if (worlds_valid && t.idx < transform_worlds.size())
		transform_worlds[t.idx] = world;
}

// SetTransformLocalMatrix / SetTransformWorldMatrix (component.cpp:236-248, :265-281). The reference does the work per call on the host;
// here calls are queued and one hgb_set_matrices launch applies the batch in call order (InverseFast(world[parent]) * world, Decompose, TRS and
// world rows written on the device).
void Scene::QueueMatrixSet(uint32_t slot, const Mat4 &m, bool is_local) {
	if (!impl->set_slots.empty() && impl->set_local != is_local)
		FlushMatrixSets();
	impl->set_local = is_local;
	impl->set_slots.push_back(slot);
	impl->set_mats.insert(impl->set_mats.end(), &m.m[0][0], &m.m[0][0] + 12);
}

void Scene::FlushMatrixSets() const {
	if (impl->set_slots.empty())
		return;
	std::vector<uint32_t> slots;
	std::vector<float> mats;
	slots.swap(impl->set_slots), mats.swap(impl->set_mats); // SyncEdits below comes back through RefreshTRSFromDevice
	Scene *self = const_cast<Scene *>(this);
	if (!self->SyncEdits(impl->resources))
		return;
	check(hgb_set_matrices(ctx, uint32_t(slots.size()), slots.data(), mats.data(), impl->set_local ? 1 : 0));
	trs_host_stale = true; // the decomposed TRS lives on the device
	worlds_valid = false;
}

void Scene::SetTransformLocalMatrix(ComponentRef ref, const Mat4 &local) {
	if (!transforms.is_valid(ref)) {
		warn("Invalid transform component");
		return;
	}
	if (ref.idx >= transform_worlds.size()) { // the reference writes transform_worlds[ref.idx] unchecked; refuse instead of corrupting
		warn("Invalid transform index");
		return;
	}
	QueueMatrixSet(ref.idx, local, true);
}

void Scene::SetTransformWorldMatrix(ComponentRef ref, const Mat4 &world) {
	if (!transforms.is_valid(ref)) {
		warn("Invalid transform component");
		return;
	}
	if (ref.idx >= transform_worlds.size()) {
		warn("Invalid transform index");
		return;
	}
	QueueMatrixSet(ref.idx, world, false);
}

Mat4 Scene::GetTransformWorldMatrix(const uint32_t transform_idx) const {
	FlushMatrixSets();
	if (!worlds_valid && ctx && !impl->topology_dirty && transform_idx < transforms.capacity()) { // one 48-byte row, not the whole mirror
		Mat4 row;
		if (hgb_get_worlds(ctx, transform_idx, 1, &row.m[0][0], 0) == 0)
			return row;
		warn(hgb_last_error());
	}
	const std::vector<Mat4> &w = GetTransformWorldMatrices();
	if (transform_idx < w.size())
		return w[transform_idx];
	warn("Invalid transform index");
	return Mat4::Identity;
}

Mat4 Scene::GetPreviousTransformWorldMatrix(const uint32_t transform_idx) const {
	const std::vector<Mat4> &p = GetPreviousTransformWorldMatrices();
	if (transform_idx < p.size())
		return p[transform_idx];
	warn("Invalid transform index");
	return GetTransformWorldMatrix(transform_idx);
}

// ---- objects -----------------------------------------------------------------------------------------------------------------
// ---- cameras (component.cpp:306-499, scene.cpp:240-263) -------------------------------------------------------------------------------
bool Camera::IsValid() const { return scene && scene->IsValidCameraRef(ref); }
float Camera::GetZNear() const { return scene ? scene->GetCameraZNear(ref) : 0.01f; }
void Camera::SetZNear(float v) { if (scene) scene->SetCameraZNear(ref, v); }
float Camera::GetZFar() const { return scene ? scene->GetCameraZFar(ref) : 1000.f; }
void Camera::SetZFar(float v) { if (scene) scene->SetCameraZFar(ref, v); }
CameraZRange Camera::GetZRange() const { return scene ? scene->GetCameraZRange(ref) : CameraZRange{}; }
void Camera::SetZRange(CameraZRange v) { if (scene) scene->SetCameraZRange(ref, v); }
float Camera::GetFov() const { return scene ? scene->GetCameraFov(ref) : 40.f / 180.f * 3.1415926535f; }
void Camera::SetFov(float v) { if (scene) scene->SetCameraFov(ref, v); }
bool Camera::GetIsOrthographic() const { return scene && scene->GetCameraIsOrthographic(ref); }
void Camera::SetIsOrthographic(bool v) { if (scene) scene->SetCameraIsOrthographic(ref, v); }
float Camera::GetSize() const { return scene ? scene->GetCameraSize(ref) : 1.f; }
void Camera::SetSize(float v) { if (scene) scene->SetCameraSize(ref, v); }

Camera Scene::CreateCamera() { return {this, cameras.add_ref(Camera_{})}; }
Camera Scene::CreateCamera(float znear, float zfar, float fov) {
	Camera_ c;
	c.zrange = {znear, zfar};
	c.fov = fov;
	return {this, cameras.add_ref(c)};
}
Camera Scene::CreateOrthographicCamera(float znear, float zfar, float size) {
	Camera_ c;
	c.zrange = {znear, zfar};
	c.ortho = true;
	c.size = size;
	return {this, cameras.add_ref(c)};
}
void Scene::DestroyCamera(ComponentRef ref) { cameras.remove_ref(ref); }

#define HG_CAM_GET(EXPR, NEUTRAL)                                                                                                                             \
	if (cameras.is_valid(ref))                                                                                                                                 \
		return cameras[ref.idx].EXPR;                                                                                                                          \
	warn("Invalid camera component");                                                                                                                          \
	return NEUTRAL;
#define HG_CAM_SET(EXPR)                                                                                                                                      \
	if (cameras.is_valid(ref))                                                                                                                                 \
		cameras[ref.idx].EXPR = v;                                                                                                                             \
	else                                                                                                                                                       \
		warn("Invalid camera component");
float Scene::GetCameraZNear(ComponentRef ref) const { HG_CAM_GET(zrange.znear, 0.01f) }
void Scene::SetCameraZNear(ComponentRef ref, float v) { HG_CAM_SET(zrange.znear) }
float Scene::GetCameraZFar(ComponentRef ref) const { HG_CAM_GET(zrange.zfar, 1000.f) }
void Scene::SetCameraZFar(ComponentRef ref, float v) { HG_CAM_SET(zrange.zfar) }
CameraZRange Scene::GetCameraZRange(ComponentRef ref) const { HG_CAM_GET(zrange, CameraZRange{}) }
void Scene::SetCameraZRange(ComponentRef ref, CameraZRange v) { HG_CAM_SET(zrange) }
float Scene::GetCameraFov(ComponentRef ref) const { HG_CAM_GET(fov, 40.f / 180.f * 3.1415926535f) }
void Scene::SetCameraFov(ComponentRef ref, float v) { HG_CAM_SET(fov) }
float Scene::GetCameraSize(ComponentRef ref) const { HG_CAM_GET(size, 1.f) }
void Scene::SetCameraSize(ComponentRef ref, float v) { HG_CAM_SET(size) }
bool Scene::GetCameraIsOrthographic(ComponentRef ref) const { HG_CAM_GET(ortho, false) }
void Scene::SetCameraIsOrthographic(ComponentRef ref, const bool &v) { HG_CAM_SET(ortho) }
Vec2 Scene::GetCameraCenterOffset(ComponentRef ref) const { HG_CAM_GET(center_offset, Vec2()) }
void Scene::SetCameraCenterOffset(ComponentRef ref, const Vec2 &v) { HG_CAM_SET(center_offset) }
#undef HG_CAM_GET
#undef HG_CAM_SET

ComponentRef Scene::GetNodeCameraRef(NodeRef ref) const {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return InvalidComponentRef;
	}
	return nodes[ref.idx].camera;
}
void Scene::SetNodeCamera(NodeRef ref, ComponentRef cref) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return;
	}
	nodes[ref.idx].camera = cref;
	impl->anims_dirty = true; // a playing "Camera.Fov" track may now have (or have lost) its target
}

// scene.cpp:240-261: the camera node's world matrix is the one the device computed this frame (one 48-byte row comes back through the
// lazily refreshed host mirror); projection and planes are built on the host with the reference's operation order (hg_math.cpp)
ViewState Scene::ComputeCameraViewState(NodeRef ref, const Vec2 &aspect_ratio) const {
	ViewState neutral{}; // the reference returns a value-initialised ViewState: every float zero (its Mat4 has no default initialiser)
	neutral.view = Mat4::Zero;
	neutral.frustum.fill(Vec4(0, 0, 0, 0));
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return neutral;
	}
	const Node_ &n = nodes[ref.idx];
	if (!transforms.is_valid(n.transform)) {
		warn("Invalid node transform");
		return neutral;
	}
	if (!cameras.is_valid(n.camera)) {
		warn("Invalid node camera");
		return neutral;
	}
	const Camera_ &cam = cameras[n.camera.idx];
	const Mat4 world = GetTransformWorldMatrix(n.transform.idx);
	return cam.ortho ? ComputeOrthographicViewState(world, cam.size, cam.zrange.znear, cam.zrange.zfar, aspect_ratio)
					 : ComputePerspectiveViewState(world, cam.fov, cam.zrange.znear, cam.zrange.zfar, aspect_ratio, Vec2(), cam.center_offset);
}

Object Scene::CreateObject() { return {this, objects.add_ref(Object_{})}; }

Object Scene::CreateObject(const ModelRef &model, std::vector<Material> materials) {
	Object_ o;
	o.model = model;
	o.materials = std::move(materials);
	return {this, objects.add_ref(std::move(o))};
}

void Scene::DestroyObject(ComponentRef ref) {
	if (objects.is_valid(ref)) {
		objects.remove_ref(ref);
		impl->touch_entries_all();
	}
}

#define HG_OBJ(RET)                                                                                                                                            \
	if (!objects.is_valid(ref)) {                                                                                                                              \
		warn("Invalid object component");                                                                                                                      \
		return RET;                                                                                                                                            \
	}
void Scene::SetObjectModel(ComponentRef ref, const ModelRef &model) {
	HG_OBJ()
	objects[ref.idx].model = model;
	impl->touch_entries_all();
}
ModelRef Scene::GetObjectModel(ComponentRef ref) const {
	HG_OBJ(InvalidModelRef)
	return objects[ref.idx].model;
}
void Scene::SetObjectMaterialCount(ComponentRef ref, size_t count) {
	HG_OBJ()
	objects[ref.idx].materials.resize(count);
	impl->touch_entries_all();
}
void Scene::SetObjectMaterial(ComponentRef ref, size_t slot, Material mat) {
	HG_OBJ()
	if (slot >= objects[ref.idx].materials.size()) {
		warn("Invalid material slot");
		return;
	}
	objects[ref.idx].materials[slot] = mat;
	impl->touch_entries_all();
}
size_t Scene::GetObjectBoneCount(ComponentRef ref) const {
	HG_OBJ(0)
	return objects[ref.idx].bones.size();
}
void Scene::SetObjectBoneCount(ComponentRef ref, size_t count) {
	HG_OBJ()
	objects[ref.idx].bones.resize(count);
	impl->touch_entries_all();
}
bool Scene::SetObjectBone(ComponentRef ref, size_t idx, NodeRef bone_node) {
	HG_OBJ(false)
	if (idx >= objects[ref.idx].bones.size()) {
		warn("Invalid bone index");
		return false;
	}
	objects[ref.idx].bones[idx] = bone_node;
	impl->touch_entries_all();
	return true;
}
NodeRef Scene::GetObjectBone(ComponentRef ref, size_t idx) const {
	HG_OBJ(InvalidNodeRef)
	return idx < objects[ref.idx].bones.size() ? objects[ref.idx].bones[idx] : InvalidNodeRef;
}
ComponentRef Scene::GetNodeObjectRef(NodeRef ref) const {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return InvalidComponentRef;
	}
	return nodes[ref.idx].object;
}
void Scene::SetNodeObject(NodeRef ref, ComponentRef cref) {
	if (!nodes.is_valid(ref)) {
		warn("Invalid node");
		return;
	}
	nodes[ref.idx].object = cref;
	impl->touch_entries(ref.idx);
}

// ---- animation ---------------------------------------------------------------------------------------------------------------
AnimRef Scene::AddAnim(Anim anim) { return anims.add_ref(std::move(anim)); }
void Scene::DestroyAnim(AnimRef ref) {
	anims.remove_ref(ref);
	impl->anims_dirty = true;
}
Anim *Scene::GetAnim(AnimRef ref) {
	impl->anims_dirty = true; // the caller may edit the tracks through the pointer
	return anims.is_valid(ref) ? &anims[ref.idx] : nullptr;
}
const Anim *Scene::GetAnim(AnimRef ref) const { return anims.is_valid(ref) ? &anims[ref.idx] : nullptr; }
SceneAnimRef Scene::AddSceneAnim(SceneAnim anim) { return scene_anims.add_ref(std::move(anim)); }
void Scene::DestroySceneAnim(SceneAnimRef ref) {
	scene_anims.remove_ref(ref);
	impl->anims_dirty = true;
}
SceneAnim *Scene::GetSceneAnim(SceneAnimRef ref) { return scene_anims.is_valid(ref) ? &scene_anims[ref.idx] : nullptr; }

ScenePlayAnimRef Scene::PlayAnim(SceneAnimRef ref, AnimLoopMode loop_mode, Easing, time_ns t_start, time_ns t_end, bool paused, float t_scale) {
	if (!scene_anims.is_valid(ref)) {
		warn("Invalid scene animation reference");
		return InvalidScenePlayAnimRef;
	}
	ScenePlayAnim_ p;
	p.scene_anim = ref;
	p.t_start = t_start != UnspecifiedAnimTime ? t_start : scene_anims[ref.idx].t_start;
	p.t_end = t_end != UnspecifiedAnimTime ? t_end : scene_anims[ref.idx].t_end;
	p.t = p.t_start;
	p.t_scale = int8_t(t_scale * 16.f); // 4-bit fixed point, 16 == 1x (scene.cpp:2371)
	p.loop_mode = loop_mode;
	p.paused = paused;
	p.play_slot = 0;
	impl->anims_dirty = true;
	return play_anims.add_ref(p);
}

bool Scene::IsPlaying(ScenePlayAnimRef ref) const { return play_anims.is_valid(ref); }
void Scene::StopAnim(ScenePlayAnimRef ref) {
	play_anims.remove_ref(ref);
	impl->anims_dirty = true;
}
void Scene::StopAllAnims() {
	play_anims.clear();
	impl->anims_dirty = true;
}

// Clock half of UpdatePlayingAnims (scene.cpp:2380-2431): exact int64 arithmetic on the host, O(#playing anims). The evaluation half
// (EvaluateBoundAnim -> Evaluate*) is the device's k_sample_tracks, fed with the times computed here.
// EvaluateHermite for a float track (animation.h:155-169 with GetIntervalKeys :109-128 and HermiteInterpolate, foundation/math.h:137-149):
// same interval rule (first key later than t), same operation order, time ratio through float(double(ns) / 1e9) like time_to_sec_f
static bool evaluate_hermite(const AnimTrackHermiteT<float> &track, time_ns t, float &v) {
	const int n = int(track.keys.size());
	if (n == 0)
		return false;
	int i = 0;
	while (i < n && !(track.keys[i].t > t))
		++i;
	if (i == 0 || i == n) {
		v = track.keys[i == 0 ? 0 : n - 1].v;
		return true;
	}
	const int k1 = i - 1, k2 = i, k0 = k1 > 0 ? k1 - 1 : 0, k3 = k2 + 1 < n ? k2 + 1 : n - 1;
	const float u = float(double(t - track.keys[k1].t) / 1000000000.0) / float(double(track.keys[k2].t - track.keys[k1].t) / 1000000000.0);
	const float y0 = track.keys[k0].v, y1 = track.keys[k1].v, y2 = track.keys[k2].v, y3 = track.keys[k3].v;
	const float tension = track.keys[k1].tension, bias = track.keys[k1].bias;
	const float u2 = u * u, u3 = u2 * u;
	float m0 = (y1 - y0) * (1.f + bias) * (1.f - tension) / 2.f;
	m0 += (y2 - y1) * (1.f - bias) * (1.f - tension) / 2.f;
	float m1 = (y2 - y1) * (1.f + bias) * (1.f - tension) / 2.f;
	m1 += (y3 - y2) * (1.f - bias) * (1.f - tension) / 2.f;
	const float a0 = 2.f * u3 - 3.f * u2 + 1.f, a1 = u3 - 2.f * u2 + u, a2 = u3 - u2, a3 = -2.f * u3 + 3.f * u2;
	v = y1 * a0 + m0 * a1 + m1 * a2 + y2 * a3;
	return true;
}

// Evaluate* for the single-value targets, generic over the number of float components (float, Color, Vec4): EvaluateHermite above
// component by component (the reference's operators on Color / Vec4 are component-wise, foundation/color.cpp:13-21, vector4.h:153-154)
template <typename T, int N> static bool evaluate_hermite_n(const AnimTrackHermiteT<T> &track, time_ns t, float *out) {
	AnimTrackHermiteT<float> one;
	one.keys.resize(track.keys.size());
	if (track.keys.empty())
		return false;
	for (int c = 0; c < N; ++c) {
		for (size_t k = 0; k < track.keys.size(); ++k) {
			const float *v = reinterpret_cast<const float *>(&track.keys[k].v);
			one.keys[k].t = track.keys[k].t, one.keys[k].v = v[c], one.keys[k].tension = track.keys[k].tension, one.keys[k].bias = track.keys[k].bias;
		}
		evaluate_hermite(one, t, out[c]);
	}
	return true;
}

// EvaluateStep (animation.h:130-139) over GetIntervalKeys (:109-128): the key at or before t, the first key before the track starts
template <typename Key> static int step_key(const std::vector<Key> &keys, time_ns t) {
	int i = 0;
	const int n = int(keys.size());
	while (i < n && !(keys[i].t > t))
		++i;
	return i > 0 ? i - 1 : 0;
}

// clock of a nested instance anim (scene.cpp:2243-2264), exact integer arithmetic
static time_ns nested_time(const AnimKeyT<InstanceAnimKey> &key, time_ns t, const SceneAnim *sub) {
	time_ns sub_t = ((t - key.t) * time_ns(key.v.t_scale * 256.f)) / 256;
	if ((key.v.t_scale < 0.f || key.v.loop_mode == ALM_Loop) && sub) {
		if (key.v.t_scale < 0.f)
			sub_t += sub->t_end;
		if (key.v.loop_mode == ALM_Loop && sub->t_end > sub->t_start) {
			if (key.v.t_scale >= 0) {
				while (sub_t >= sub->t_end)
					sub_t -= sub->t_end - sub->t_start;
			} else {
				while (sub_t <= sub->t_start)
					sub_t += sub->t_end - sub->t_start;
			}
		}
	}
	return sub_t;
}
// the key of an instance_anim_track in force at t: the last key with key.t <= t, -1 before the first (scene.cpp:2222-2225)
static int instance_key(const Anim &anim, time_ns t) {
	int kf = int(anim.instance_anim_track.keys.size()) - 1;
	for (; kf >= 0; --kf)
		if (t >= anim.instance_anim_track.keys[kf].t)
			break;
	return kf;
}

void Scene::EvaluateHostTracks() {
	for (const Impl::HostTrack &ht : impl->host_tracks) {
		if (ht.clock >= impl->play_t.size() || impl->play_t[ht.clock] == HGB_NOT_PLAYING || !anims.is_valid(ht.anim))
			continue;
		const time_ns t = impl->play_t[ht.clock];
		const Anim &anim = anims[ht.anim.idx];
		switch (ht.target) {
		case Impl::HT_CameraFov:
			if (cameras.is_valid(ht.comp) && size_t(ht.track) < anim.float_tracks.size())
				evaluate_hermite(anim.float_tracks[ht.track], t, cameras[ht.comp.idx].fov);
			break;
		case Impl::HT_LightDiffuseIntensity:
			if (lights.is_valid(ht.comp) && size_t(ht.track) < anim.float_tracks.size())
				evaluate_hermite(anim.float_tracks[ht.track], t, lights[ht.comp.idx].diffuse_intensity);
			break;
		case Impl::HT_LightSpecularIntensity:
			if (lights.is_valid(ht.comp) && size_t(ht.track) < anim.float_tracks.size())
				evaluate_hermite(anim.float_tracks[ht.track], t, lights[ht.comp.idx].specular_intensity);
			break;
		case Impl::HT_LightDiffuse:
			if (lights.is_valid(ht.comp) && size_t(ht.track) < anim.color_tracks.size())
				evaluate_hermite_n<Color, 4>(anim.color_tracks[ht.track], t, &lights[ht.comp.idx].diffuse.r);
			break;
		case Impl::HT_LightSpecular:
			if (lights.is_valid(ht.comp) && size_t(ht.track) < anim.color_tracks.size())
				evaluate_hermite_n<Color, 4>(anim.color_tracks[ht.track], t, &lights[ht.comp.idx].specular.r);
			break;
		case Impl::HT_MaterialValue:
			if (objects.is_valid(ht.comp) && size_t(ht.track) < anim.vec4_tracks.size()) {
				Object_ &obj = objects[ht.comp.idx];
				if (ht.mat_slot >= obj.materials.size())
					break; // invalid material slot
				auto i = obj.materials[ht.mat_slot].values.find(ht.mat_value);
				if (i == obj.materials[ht.mat_slot].values.end())
					break; // invalid material value name
				float v[4];
				if (evaluate_hermite_n<Vec4, 4>(anim.vec4_tracks[ht.track], t, v))
					i->second.value = {v[0], v[1], v[2], v[3]};
			}
			break;
		case Impl::HT_EnableInstanceHost: // "Enable" on a node that hosts an instance: the switch propagates to the instantiated nodes
			if (nodes.is_valid(ht.node) && size_t(ht.track) < anim.bool_tracks.size() && !anim.bool_tracks[ht.track].keys.empty()) {
				const auto &keys = anim.bool_tracks[ht.track].keys;
				const bool enable = keys[step_key(keys, t)].v;
				if (enable != IsNodeItselfEnabled(ht.node))
					enable ? EnableNode(ht.node) : DisableNode(ht.node);
			}
			break;
		}
	}
}

void Scene::UpdatePlayingAnims(time_ns dt) {
	// clocks first (scene.cpp:2386-2413): they do not depend on the bindings
	std::vector<uint32_t> finished;
	for (uint32_t i = play_anims.first(); i != play_anims.invalid_idx; i = play_anims.next(i)) {
		ScenePlayAnim_ &p = play_anims[i];
		if (!p.paused)
			p.t += (dt * p.t_scale) >> 4;
		const time_ns span = p.t_end - p.t_start;
		if (p.loop_mode == ALM_Loop) {
			if (p.t_scale >= 0) {
				while (p.t >= p.t_end && span > 0)
					p.t -= span;
			} else {
				while (p.t <= p.t_start && span > 0)
					p.t += span;
			}
		} else if (p.loop_mode == ALM_Once) {
			if (p.t_scale >= 0 ? p.t >= p.t_end : p.t <= p.t_start) {
				p.t = p.t_scale >= 0 ? p.t_end : p.t_start;
				finished.push_back(i); // evaluated one last time this frame, then removed
			}
		}
	}
	if (impl->anims_dirty)
		SyncEdits(impl->resources); // (re)binds tracks and assigns play slots

	// nested instance anims: the key in force may have moved on -> the tracks of another scene anim must be bound
	std::vector<time_ns> nt;
	std::vector<uint8_t> active;
	auto walk_nested = [&]() -> bool {
		bool changed = false;
		nt.assign(impl->nested.size(), 0);
		active.assign(impl->nested.size(), 0);
		for (size_t i = 0; i < impl->nested.size(); ++i) {
			Impl::Nested &r = impl->nested[i];
			time_ns t_parent;
			if (r.parent < 0) {
				if (!play_anims.is_used(r.play_idx))
					continue;
				t_parent = play_anims[r.play_idx].t;
			} else {
				if (!active[r.parent] || !scene_anims.is_valid(impl->nested[r.parent].sub))
					continue;
				t_parent = nt[r.parent];
			}
			if (!anims.is_valid(r.anim))
				continue;
			const Anim &anim = anims[r.anim.idx];
			const int kf = instance_key(anim, t_parent);
			if (kf != r.kf)
				changed = true;
			active[i] = 1;
			if (kf >= 0 && kf == r.kf && scene_anims.is_valid(r.sub))
				nt[i] = nested_time(anim.instance_anim_track.keys[kf], t_parent, &scene_anims[r.sub.idx]);
		}
		return changed;
	};
	if (walk_nested()) {
		impl->anims_dirty = true;
		SyncEdits(impl->resources);
		walk_nested();
	}

	std::fill(impl->play_t.begin(), impl->play_t.end(), HGB_NOT_PLAYING);
	for (uint32_t i = play_anims.first(); i != play_anims.invalid_idx; i = play_anims.next(i))
		if (play_anims[i].play_slot < impl->play_t.size())
			impl->play_t[play_anims[i].play_slot] = play_anims[i].t;
	for (size_t i = 0; i < impl->nested.size(); ++i)
		if (active[i] && impl->nested[i].kf >= 0 && scene_anims.is_valid(impl->nested[i].sub) && impl->nested[i].play_slot < impl->play_t.size())
			impl->play_t[impl->nested[i].play_slot] = nt[i];

	EvaluateHostTracks(); // finished plays are evaluated one last time, like every other target
	for (uint32_t i : finished)
		play_anims.remove(i);
	if (!finished.empty())
		impl->rebind_after_frame = true; // re-bound (without the finished plays) once this frame's sampling has been enqueued
	if (!impl->play_t.empty())
		trs_host_stale = true;
}

// ---- edit flush: host scene graph -> device context ------------------------------------------------------------------------------
// bounds table: one MinMax per (model, list) of the resource cache, indexed through `base` of the model
static void build_bounds(const PipelineResources &res, std::vector<float> &bounds6, std::vector<uint32_t> &base_of_model) {
	bounds6.clear();
	base_of_model.assign(0x10000, 0xffffffffu);
	for (uint32_t m = res.models.first(); m != 0xffffffffu; m = res.models.next(m)) {
		const Model &mdl = res.models.Get_unsafe_(uint16_t(m));
		base_of_model[m] = uint32_t(bounds6.size() / 6);
		for (const MinMax &b : mdl.bounds) {
			const float v[6] = {b.mn.x, b.mn.y, b.mn.z, b.mx.x, b.mx.y, b.mx.z};
			bounds6.insert(bounds6.end(), v, v + 6);
		}
	}
}

bool Scene::SyncEdits(const PipelineResources *resources) {
	if (!EnsureContext())
		return false;
	FlushMatrixSets(); // queued SetTransformWorld/LocalMatrix calls come before whatever was edited after them
	RefreshFlagsFromDevice(); // node flags are about to be read (entry tables) and the set of tracked nodes may change
	const uint32_t n = uint32_t(transforms.capacity());

	if (impl->topology_dirty && impl->topology_reparent_only && !impl->reparented.empty() && impl->device_slots == n && impl->fresh_slots.empty()) {
		// only SetTransformParent calls since the last sync: the incremental path (parent rows only when no level changes)
		std::vector<uint32_t> slots, parents;
		for (uint32_t i : impl->reparented) {
			if (!transforms.is_used(i))
				continue;
			uint32_t par = HGB_ROOT;
			const NodeRef pn = transforms[i].parent;
			if (nodes.is_valid(pn)) {
				const ComponentRef pt = nodes[pn.idx].transform;
				if (transforms.is_valid(pt) && pt.idx != i)
					par = pt.idx;
			}
			slots.push_back(i), parents.push_back(par);
		}
		int rebuilt = 0;
		if (slots.empty() || hgb_set_parents(ctx, uint32_t(slots.size()), slots.data(), parents.data(), &rebuilt) == 0) {
			impl->topology_dirty = false;
			if (rebuilt) {
				RefreshTRSFromDevice();
				impl->trs_all_dirty = true;
			}
		} else {
			warn(hgb_last_error());
		}
	}
	impl->topology_reparent_only = false;
	impl->reparented.clear();
	if (impl->topology_dirty && impl->topology_append_only && impl->device_slots > 0 && !impl->fresh_slots.empty()) {
		// only transforms created (and parented) since the last sync: they take free positions of the device's append region
		uint32_t cap = 0;
		if (hgb_slot_capacity(ctx, &cap) == 0 && n <= cap) {
			std::vector<uint32_t> slots, parents;
			std::vector<float> trs;
			for (uint32_t i : impl->fresh_slots) {
				if (!transforms.is_used(i))
					continue;
				uint32_t par = HGB_ROOT;
				const NodeRef pn = transforms[i].parent;
				if (nodes.is_valid(pn)) {
					const ComponentRef pt = nodes[pn.idx].transform;
					if (transforms.is_valid(pt) && pt.idx != i)
						par = pt.idx;
				}
				slots.push_back(i), parents.push_back(par);
				const float *t9 = &transforms[i].TRS.pos.x;
				trs.insert(trs.end(), t9, t9 + 9);
			}
			// a parent created in the same batch must precede its children: hgb_append_nodes resolves parents in order
			std::vector<uint32_t> order(slots.size());
			for (size_t k = 0; k < order.size(); ++k)
				order[k] = uint32_t(k);
			{
				std::unordered_map<uint32_t, uint32_t> at;
				for (size_t k = 0; k < slots.size(); ++k)
					at[slots[k]] = uint32_t(k);
				std::vector<uint8_t> done(slots.size(), 0);
				std::vector<uint32_t> sorted;
				sorted.reserve(slots.size());
				for (size_t k = 0; k < slots.size(); ++k) { // emit the chain of not-yet-emitted fresh ancestors, root first
					std::vector<uint32_t> chain;
					uint32_t cur = uint32_t(k);
					while (!done[cur] && chain.size() <= slots.size()) {
						chain.push_back(cur);
						done[cur] = 1;
						const auto it = at.find(parents[cur]);
						if (it == at.end())
							break;
						cur = it->second;
					}
					sorted.insert(sorted.end(), chain.rbegin(), chain.rend());
				}
				order.swap(sorted);
			}
			std::vector<uint32_t> s2(slots.size()), p2(slots.size());
			std::vector<float> t2(trs.size());
			for (size_t k = 0; k < order.size(); ++k) {
				s2[k] = slots[order[k]], p2[k] = parents[order[k]];
				memcpy(&t2[k * 9], &trs[size_t(order[k]) * 9], 36);
			}
			int rebuilt = 0;
			if (s2.empty() || (hgb_append_nodes(ctx, uint32_t(s2.size()), s2.data(), p2.data(), t2.data(), &rebuilt) == 0 && !rebuilt)) {
				impl->topology_dirty = false;
				impl->device_slots = n;
				impl->fresh_slots.clear();
			}
		}
	}
	impl->topology_append_only = false;
	if (impl->topology_dirty) {
		RefreshTRSFromDevice(); // sampled values must survive the re-layout: they are re-uploaded below
		std::vector<uint32_t> parent(n, HGB_ROOT);
		std::vector<uint8_t> used(n, 0);
		for (uint32_t i = transforms.first(); i != transforms.invalid_idx; i = transforms.next(i)) {
			used[i] = 1;
			const NodeRef pn = transforms[i].parent; // parent is a NODE: hop node -> its transform (scene.cpp:268)
			if (nodes.is_valid(pn)) {
				const ComponentRef pt = nodes[pn.idx].transform;
				if (transforms.is_valid(pt) && pt.idx != i)
					parent[i] = pt.idx;
			}
		}
		if (hgb_set_topology(ctx, n, parent.data(), used.data())) {
			warn(hgb_last_error());
			return false;
		}
		impl->topology_dirty = false;
		impl->trs_all_dirty = true;
		impl->device_slots = n;
		impl->fresh_slots.clear();
	}

	if (impl->trs_all_dirty && n) {
		impl->staging.assign(size_t(n) * 9, 0.f);
		for (uint32_t i = transforms.first(); i != transforms.invalid_idx; i = transforms.next(i))
			memcpy(impl->staging.data() + size_t(i) * 9, &transforms[i].TRS, 36);
		check(hgb_upload_trs(ctx, 0, n, impl->staging.data()));
	} else if (!impl->dirty_trs.empty()) {
		std::vector<uint32_t> &d = impl->dirty_trs;
		std::sort(d.begin(), d.end());
		d.erase(std::unique(d.begin(), d.end()), d.end());
		for (size_t a = 0; a < d.size();) { // contiguous runs of dirty slots -> one upload each
			size_t b = a + 1;
			while (b < d.size() && d[b] == d[b - 1] + 1)
				++b;
			impl->staging.resize((b - a) * 9);
			for (size_t k = a; k < b; ++k)
				if (transforms.is_used(d[k]))
					memcpy(impl->staging.data() + (k - a) * 9, &transforms[d[k]].TRS, 36);
			check(hgb_upload_trs(ctx, d[a], uint32_t(b - a), impl->staging.data()));
			a = b;
		}
	}
	impl->trs_all_dirty = false;
	impl->dirty_trs.clear();

	// bound tracks (BindNodeAnim scene.cpp:2070-2152: the LAST track with a matching target name wins; AF_UseQuaternionForRotation
	// selects the quaternion "Rotation" track and ignores the Euler one, EvaluateBoundAnim :2168-2177), flattened in the reference's
	// evaluation order: playing anims by index, node anims in order, Position / Rotation / Scale
	if (impl->anims_dirty) {
		std::vector<int64_t> v3_t, q_t;
		std::vector<float> v3_val, v3_tension, v3_bias, q_val;
		std::vector<hgb_track> tracks, en_tracks;
		std::vector<int64_t> en_t;
		std::vector<uint8_t> en_v;
		impl->enable_nodes.clear();
		impl->host_tracks.clear();
		impl->nested.clear();
		uint32_t n_play = 0;
		for (uint32_t pi = play_anims.first(); pi != play_anims.invalid_idx; pi = play_anims.next(pi))
			play_anims[pi].play_slot = n_play++;
		if (n_play >= 65535)
			warn("Too many playing anims for one device context (65535)");

		// one scene anim evaluated at the clock of `slot` (a play, or a nested instance anim): its node anims in order, for each the
		// host-side single values, the "Enable" switch, Position / Rotation / Scale, then the instantiated scene's anim its
		// instance_anim_track selects at `t_now` (recursion, like EvaluateBoundAnim scene.cpp:2154-2268)
		std::function<void(const SceneAnim &, uint32_t, time_ns, int, uint32_t, int)> flatten = [&](const SceneAnim &sa, uint32_t slot, time_ns t_now, int parent,
																									 uint32_t play_idx, int depth) {
			for (const NodeAnim &na : sa.node_anims) {
				if (!anims.is_valid(na.anim) || !nodes.is_valid(na.node) || slot >= 65535)
					continue;
				const Anim &anim = anims[na.anim.idx];
				const Node_ &node = nodes[na.node.idx];
				const bool hosts_instance = node_instance_view.find(na.node) != node_instance_view.end();
				int en = -1; // "Enable" bool track: the last one with that target (BindNodeAnim scene.cpp:2083-2089)
				for (size_t i = 0; i < anim.bool_tracks.size(); ++i)
					if (anim.bool_tracks[i].target == "Enable")
						en = int(i);
				const ComponentRef t = node.transform;
				if (en >= 0 && (hosts_instance || !transforms.is_valid(t))) { // the switch reaches other nodes too / has no device row: host side
					impl->host_tracks.push_back({Impl::HT_EnableInstanceHost, slot, InvalidComponentRef, na.node, na.anim, en, 0, {}});
				} else if (en >= 0) {
					const auto &keys = anim.bool_tracks[en].keys;
					en_tracks.push_back(hgb_track{t.idx, uint32_t(en_t.size()), uint32_t(keys.size()), uint16_t(slot), 4, 0});
					for (const auto &k : keys)
						en_t.push_back(k.t), en_v.push_back(k.v ? 1 : 0);
					impl->enable_nodes.push_back({t.idx, na.node.idx});
				}
				if (transforms.is_valid(t)) {
					int pos = -1, rot = -1, scl = -1, qrot = -1;
					for (size_t i = 0; i < anim.vec3_tracks.size(); ++i) {
						const std::string &tg = anim.vec3_tracks[i].target;
						if (tg == "Position") pos = int(i);
						else if (tg == "Rotation") rot = int(i);
						else if (tg == "Scale") scl = int(i);
					}
					for (size_t i = 0; i < anim.quat_tracks.size(); ++i)
						if (anim.quat_tracks[i].target == "Rotation")
							qrot = int(i);
					auto add_v3 = [&](int idx, uint8_t kind) {
						if (idx < 0)
							return;
						const auto &keys = anim.vec3_tracks[idx].keys;
						hgb_track tr{t.idx, uint32_t(v3_t.size()), uint32_t(keys.size()), uint16_t(slot), kind, 0};
						for (const auto &k : keys) {
							v3_t.push_back(k.t);
							v3_val.push_back(k.v.x), v3_val.push_back(k.v.y), v3_val.push_back(k.v.z);
							v3_tension.push_back(k.tension), v3_bias.push_back(k.bias);
						}
						tracks.push_back(tr);
					};
					add_v3(pos, 0);
					if (anim.flags & AF_UseQuaternionForRotation) {
						if (qrot >= 0) {
							const auto &keys = anim.quat_tracks[qrot].keys;
							hgb_track tr{t.idx, uint32_t(q_t.size()), uint32_t(keys.size()), uint16_t(slot), 3, 0};
							for (const auto &k : keys) {
								q_t.push_back(k.t);
								q_val.push_back(k.v.x), q_val.push_back(k.v.y), q_val.push_back(k.v.z), q_val.push_back(k.v.w);
							}
							tracks.push_back(tr);
						}
					} else {
						add_v3(rot, 1);
					}
					add_v3(scl, 2);
				}
				// single host-side values, in the reference's order: light, camera, material (the LAST track with a matching target wins)
				if (lights.is_valid(node.light)) {
					int dif = -1, spc = -1, dif_i = -1, spc_i = -1;
					for (size_t i = 0; i < anim.color_tracks.size(); ++i) {
						if (anim.color_tracks[i].target == "Light.Diffuse") dif = int(i);
						else if (anim.color_tracks[i].target == "Light.Specular") spc = int(i);
					}
					for (size_t i = 0; i < anim.float_tracks.size(); ++i) {
						if (anim.float_tracks[i].target == "Light.DiffuseIntensity") dif_i = int(i);
						else if (anim.float_tracks[i].target == "Light.SpecularIntensity") spc_i = int(i);
					}
					if (dif >= 0) impl->host_tracks.push_back({Impl::HT_LightDiffuse, slot, node.light, na.node, na.anim, dif, 0, {}});
					if (spc >= 0) impl->host_tracks.push_back({Impl::HT_LightSpecular, slot, node.light, na.node, na.anim, spc, 0, {}});
					if (dif_i >= 0) impl->host_tracks.push_back({Impl::HT_LightDiffuseIntensity, slot, node.light, na.node, na.anim, dif_i, 0, {}});
					if (spc_i >= 0) impl->host_tracks.push_back({Impl::HT_LightSpecularIntensity, slot, node.light, na.node, na.anim, spc_i, 0, {}});
				}
				if (cameras.is_valid(node.camera)) { // the last "Camera.Fov" float track (BindNodeAnim scene.cpp:2091-2101)
					int fov = -1;
					for (size_t i = 0; i < anim.float_tracks.size(); ++i)
						if (anim.float_tracks[i].target == "Camera.Fov")
							fov = int(i);
					if (fov >= 0)
						impl->host_tracks.push_back({Impl::HT_CameraFov, slot, node.camera, na.node, na.anim, fov, 0, {}});
				}
				if (objects.is_valid(node.object))
					for (size_t i = 0; i < anim.vec4_tracks.size(); ++i) { // "Material.<slot>.<value>" (SplitMaterialPropertyName scene.cpp:2056-2068)
						const std::string &tg = anim.vec4_tracks[i].target;
						if (tg.compare(0, 9, "Material.") != 0)
							continue;
						const size_t dot = tg.find('.', 9);
						if (dot == std::string::npos || tg.find('.', dot + 1) != std::string::npos)
							continue;
						impl->host_tracks.push_back({Impl::HT_MaterialValue, slot, node.object, na.node, na.anim, int(i), uint8_t(atoi(tg.substr(9, dot - 9).c_str())),
							tg.substr(dot + 1)});
					}
				// the instantiated scene's anim selected by the step track (scene.cpp:2217-2268)
				if (!anim.instance_anim_track.keys.empty() && hosts_instance && depth < 8) {
					const SceneView &view = node_instance_view.find(na.node)->second;
					Impl::Nested r{parent, play_idx, na.node, na.anim, instance_key(anim, t_now), InvalidSceneAnimRef, 0};
					if (r.kf >= 0)
						r.sub = view.GetSceneAnim(*this, anim.instance_anim_track.keys[r.kf].v.anim_name);
					const bool bound = scene_anims.is_valid(r.sub);
					if (bound)
						r.play_slot = n_play++;
					impl->nested.push_back(r);
					if (bound) {
						const int me = int(impl->nested.size()) - 1;
						const SceneAnim &sub = scene_anims[r.sub.idx];
						flatten(sub, r.play_slot, nested_time(anim.instance_anim_track.keys[r.kf], t_now, &sub), me, play_idx, depth + 1);
					}
				}
			}
		};
		for (uint32_t pi = play_anims.first(); pi != play_anims.invalid_idx; pi = play_anims.next(pi)) {
			const ScenePlayAnim_ &p = play_anims[pi];
			if (scene_anims.is_valid(p.scene_anim))
				flatten(scene_anims[p.scene_anim.idx], p.play_slot, p.t, -1, pi, 0);
		}
		impl->n_play_slots = n_play;
		check(hgb_set_anim_keys(ctx, uint32_t(v3_t.size()), v3_t.data(), v3_val.data(), v3_tension.data(), v3_bias.data(), uint32_t(q_t.size()), q_t.data(),
			q_val.data()));
		check(hgb_set_tracks(ctx, uint32_t(tracks.size()), tracks.data(), n_play));
		check(hgb_set_enable_tracks(ctx, uint32_t(en_tracks.size()), en_tracks.data(), uint32_t(en_t.size()), en_t.data(), en_v.data()));
		impl->enable_state_push = !impl->enable_nodes.empty();
		impl->play_t.assign(n_play, HGB_NOT_PLAYING);
		impl->anims_dirty = false;
	}

	// fused culling (extension): the device entry tables = every rigid display entry in GetModelDisplayLists order, disabled nodes
	// flagged instead of skipped, so that Enable/DisableNode only changes a flag
	const bool can_fuse = resources && !impl->frusta.empty();
	const bool append = can_fuse && impl->entries_dirty && impl->entries_append_only && impl->ent_built;
	if (can_fuse && impl->ent_built && (!impl->entries_dirty || append) && !impl->flag_nodes.empty()) {
		// Enable / DisableNode since the last sync: the entries of those nodes flip their flag where the kernels read it
		std::vector<uint32_t> &fn = impl->flag_nodes;
		std::sort(fn.begin(), fn.end());
		fn.erase(std::unique(fn.begin(), fn.end()), fn.end());
		for (uint32_t ni : fn) {
			if (!nodes.is_used(ni))
				continue;
			const uint8_t flag = (nodes[ni].flags & (NF_Disabled | NF_InstanceDisabled)) ? HGB_ENTRY_DISABLED : 0;
			for (int k = 0; k < 2; ++k) {
				const auto r = std::equal_range(impl->ent_node[k].begin(), impl->ent_node[k].end(), ni);
				const uint32_t e0 = uint32_t(r.first - impl->ent_node[k].begin()), e1 = uint32_t(r.second - impl->ent_node[k].begin());
				if (e0 == e1 || impl->ent_flags[k][e0] == flag)
					continue;
				std::fill(impl->ent_flags[k].begin() + e0, impl->ent_flags[k].begin() + e1, flag);
				check(hgb_set_entry_flags(ctx, k, e0, e1 - e0, impl->ent_flags[k].data() + e0));
			}
		}
	}
	impl->flag_nodes.clear();
	if (can_fuse && impl->entries_dirty) {
		std::vector<float> bounds6;
		std::vector<uint32_t> base_of_model;
		build_bounds(*resources, bounds6, base_of_model);
		std::vector<uint32_t> mtx[2], bidx[2];
		size_t old_n[2] = {0, 0};
		if (!append)
			for (int k = 0; k < 2; ++k)
				impl->ent[k].clear(), impl->ent_flags[k].clear(), impl->ent_node[k].clear();
		for (int k = 0; k < 2; ++k)
			old_n[k] = impl->ent[k].size();
		// an append walks only the nodes behind the last one the tables cover (touch_entries saw no edit below it)
		uint32_t ni = nodes.first();
		if (append) {
			ni = impl->ent_nodes_end;
			while (ni < nodes.capacity() && !nodes.is_used(ni))
				++ni;
			if (ni >= nodes.capacity())
				ni = nodes.invalid_idx;
		}
		uint32_t last_node = append && impl->ent_nodes_end ? impl->ent_nodes_end - 1 : 0;
		for (; ni != nodes.invalid_idx; ni = nodes.next(ni)) {
			last_node = ni;
			const Node_ &node = nodes[ni];
			if (!transforms.is_valid(node.transform) || !objects.is_valid(node.object))
				continue;
			const Object_ &obj = objects[node.object.idx];
			const uint16_t mdl_idx = resources->models.GetValidatedRefIndex(obj.model);
			if (mdl_idx == 0xffff)
				continue;
			const Model &mdl = resources->models.Get_unsafe_(mdl_idx);
			if (!obj.bones.empty() && obj.bones.size() == mdl.bind_pose.size())
				continue; // skinned lists are never culled (scene_forward_pipeline.cpp:560-562)
			const uint8_t flag = (node.flags & (NF_Disabled | NF_InstanceDisabled)) ? HGB_ENTRY_DISABLED : 0;
			for (size_t i = 0; i < mdl.lists.size() && i < mdl.bounds.size(); ++i) {
				const uint16_t mat_idx = mdl.mats[i];
				if (mat_idx >= obj.materials.size())
					continue;
				const int k = GetMaterialBlendMode(obj.materials[mat_idx]) != BM_Opaque ? 1 : 0;
				impl->ent[k].push_back({&obj.materials[mat_idx], node.transform.idx, mdl_idx, uint16_t(i)});
				impl->ent_flags[k].push_back(flag);
				impl->ent_node[k].push_back(ni);
				mtx[k].push_back(node.transform.idx);
				bidx[k].push_back(base_of_model[mdl_idx] + uint32_t(i));
			}
		}
		static_assert(sizeof(ModelDisplayList) == 16, "the device keeps one 16-byte record per entry");
		if (!append || bounds6.size() / 6 != impl->ent_bounds_count) {
			check(hgb_set_bounds(ctx, uint32_t(bounds6.size() / 6), bounds6.data()));
			impl->ent_bounds_count = bounds6.size() / 6;
		}
		for (int k = 0; k < 2; ++k) {
			if (append) {
				const uint32_t cnt = uint32_t(mtx[k].size());
				if (!cnt)
					continue;
				int rebuilt = 0;
				check(hgb_append_display_entries(ctx, k, cnt, mtx[k].data(), bidx[k].data(), impl->ent_flags[k].data() + old_n[k], nullptr, &rebuilt));
				if (hgb_set_entry_records(ctx, k, uint32_t(old_n[k]), cnt, impl->ent[k].data() + old_n[k])) // the table outgrew its slack: all of them again
					check(hgb_set_entry_records(ctx, k, 0, uint32_t(impl->ent[k].size()), impl->ent[k].data()));
				continue;
			}
			check(hgb_set_display_entries(ctx, k, uint32_t(mtx[k].size()), mtx[k].data(), bidx[k].data(), impl->ent_flags[k].data(), nullptr));
			if (!impl->ent[k].empty()) // the display lists themselves: survivors come back as these records (GetVisibleModelDisplayLists)
				check(hgb_set_entry_records(ctx, k, 0, uint32_t(impl->ent[k].size()), impl->ent[k].data()));
		}
		if (!append && impl->append_slack) // later appends must not re-allocate (and copy) the whole table
			for (int k = 0; k < 2; ++k) {
				const size_t want = impl->ent[k].size() + size_t(impl->append_slack);
				impl->ent[k].reserve(want), impl->ent_flags[k].reserve(want), impl->ent_node[k].reserve(want);
			}
		impl->ent_nodes_end = last_node + 1;
		impl->ent_built = true;
		impl->entries_dirty = false;
	}
	impl->entries_append_only = false;
	++impl->epoch; // nothing created before this sync counts as fresh any more
	if (impl->enable_state_push && !impl->enable_nodes.empty()) { // the host's flags are the truth right now: tracked nodes take them
		std::vector<uint32_t> slots;
		std::vector<uint8_t> dis;
		for (const auto &sn : impl->enable_nodes)
			if (nodes.is_used(sn.second)) {
				slots.push_back(sn.first);
				dis.push_back((nodes[sn.second].flags & (NF_Disabled | NF_InstanceDisabled)) ? 1 : 0);
			}
		check(hgb_set_enable_state(ctx, uint32_t(slots.size()), slots.data(), dis.data()));
	}
	impl->enable_state_push = false;
	if (impl->views_dirty && !impl->frusta.empty()) {
		check(hgb_set_views(ctx, uint32_t(impl->frusta.size()), &impl->frusta[0][0].x));
		impl->views_dirty = false;
	}
	return true;
}

void Scene::SetCullingFrustums(const std::vector<Frustum> &frustums, const PipelineResources &resources) {
	if (frustums.size() > HGB_MAX_VIEWS) {
		warn("Too many culling frustums (8 at most)");
		return;
	}
	if (impl->resources != &resources || impl->frusta.empty() != frustums.empty())
		impl->touch_entries_all();
	impl->resources = &resources;
	impl->frusta = frustums;
	impl->views_dirty = true;
}

// where the time of GetVisibleModelDisplayLists goes (seconds, accumulated): [0] the device round trip (frame tail + gather + copy to
// pinned memory), [1] sizing the caller's vector, [2] the copy into it. Read and cleared by the flat binding's timing leg.
double g_visible_breakdown[3] = {0, 0, 0};
static inline double now_sec() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

bool Scene::GetVisibleModelDisplayLists(uint32_t view, bool transparent, std::vector<ModelDisplayList> &out, const PipelineResources &) {
	const int k = transparent ? 1 : 0;
	const double t0 = now_sec();
	if (!ctx || impl->frusta.empty() || view >= impl->frusta.size() || impl->entries_dirty) {
		out.clear();
		warn("GetVisibleModelDisplayLists: no fused culling result for this view (SetCullingFrustums + Update first)");
		return false;
	}
	// the survivors' ModelDisplayList records, gathered on the device (a sparse walk through a table of 16 bytes per entry is what a GPU
	// is good at) and landed contiguously in a pinned buffer: count and records in one round trip
	const void *rec = nullptr;
	uint32_t n = 0;
	constexpr uint32_t kChunkRecords = 1u << 16; // 1 MB
	if (hgb_get_visible_records_begin(ctx, k, view, kChunkRecords, &rec, &n)) {
		warn(hgb_last_error());
		return false;
	}
	// The count is known, the records are still arriving chunk by chunk. ModelDisplayList is trivially copyable: size the vector (a
	// caller that reuses its vector pays no re-initialisation when the count barely moves) and let the pool copy every chunk the
	// moment it has landed — the copy into the caller's vector hides behind the link
	static_assert(std::is_trivially_copyable<ModelDisplayList>::value, "copied with memcpy");
	const double t1 = now_sec();
	if (out.capacity() < n) // a list that grows by a few entries per frame must not re-allocate (and page in) 26 MB every time it does
		out.reserve(size_t(n) + n / 8 + 1024);
	out.resize(n);
	const double t2 = now_sec();
	const uint8_t *src = static_cast<const uint8_t *>(rec);
	uint8_t *dst = reinterpret_cast<uint8_t *>(out.data());
	const size_t bytes = size_t(n) * sizeof(ModelDisplayList), chunk = size_t(kChunkRecords) * sizeof(ModelDisplayList);
	const uint32_t n_chunks = uint32_t((bytes + chunk - 1) / chunk);
	hgb_ctx *dev = ctx;
	std::atomic<int> failed{0};
	GatherPool::get().run(n_chunks, [=, &failed](uint32_t c) {
		if (hgb_visible_records_wait(dev, k, c))
			failed.store(1);
		const size_t b = size_t(c) * chunk;
		memcpy(dst + b, src + b, std::min(chunk, bytes - b));
	});
	if (failed.load()) {
		warn("GetVisibleModelDisplayLists: the device copy failed");
		out.clear();
		return false;
	}
	const double t3 = now_sec();
	g_visible_breakdown[0] += t1 - t0, g_visible_breakdown[1] += t2 - t1, g_visible_breakdown[2] += t3 - t2;
	return true;
}

// ---- per-frame path ----------------------------------------------------------------------------------------------------------
// a2 (scene.cpp:291-294): the device swaps its two world buffers inside hgb_update; the host mirrors swap with them.
void Scene::StorePreviousWorldMatrices() {
	std::swap(transform_worlds, previous_transform_worlds);
	std::swap(worlds_valid, previous_valid);
	// the registry is keyed by vector ADDRESS, which std::swap does not change: both addresses keep naming this scene
}

// a3 (scene.cpp:280-284): sizes the world vectors to transforms.capacity() and opens the frame: edits are flushed to the device and
// SetNodeWorldMatrix calls from here to ComputeWorldMatrices are recorded as presets.
void Scene::ReadyWorldMatrices() {
	transform_worlds.resize(transforms.capacity());
	worlds_valid = false;
	impl->preset_slots.clear();
	impl->preset_mats.clear();
	impl->in_frame = true;
	SyncEdits(impl->resources);
}

// a4 (scene.cpp:263-289) — on the device: k_sample_tracks + k_update_level per hierarchy level (+ fused visibility / compaction)
void Scene::ComputeWorldMatrices() {
	impl->in_frame = false;
	if (!SyncEdits(impl->resources)) // setters may have run since ReadyWorldMatrices (scripts, physics)
		return;
	if (!impl->preset_slots.empty())
		check(hgb_set_preset_world(ctx, uint32_t(impl->preset_slots.size()), impl->preset_slots.data(), impl->preset_mats.data()));
	const bool fused = impl->resources && !impl->frusta.empty();
	check(hgb_update(ctx, uint32_t(impl->play_t.size()), impl->play_t.data(), fused ? 1 : 0));
	if (!impl->enable_nodes.empty())
		for (const int64_t t : impl->play_t)
			if (t != HGB_NOT_PLAYING) {
				impl->flags_stale = true; // Enable tracks were evaluated on the device
				break;
			}
	std::fill(impl->play_t.begin(), impl->play_t.end(), HGB_NOT_PLAYING); // a clock drives the sampling of ONE frame
	if (impl->rebind_after_frame)
		impl->anims_dirty = true, impl->rebind_after_frame = false;
	impl->preset_slots.clear();
	impl->preset_mats.clear();
	worlds_valid = false;
}

// a5 (scene.cpp:296-304): the reference resizes the previous vectors; its copy loop is a no-op (`==` at :303)
void Scene::FixupPreviousWorldMatrices() { previous_transform_worlds.resize(transform_worlds.size()); }

void Scene::Update(time_ns dt) { // scene.cpp:307-315
	StorePreviousWorldMatrices();
	ReadyWorldMatrices();
	UpdatePlayingAnims(dt);
	ComputeWorldMatrices();
	FixupPreviousWorldMatrices();
}

static void download_worlds(hgb_ctx *ctx, std::vector<Mat4> &dst, size_t capacity, bool previous) {
	dst.resize(capacity);
	if (ctx && capacity)
		check(hgb_get_worlds(ctx, 0, uint32_t(capacity), &dst[0].m[0][0], previous ? 1 : 0));
}

const std::vector<Mat4> &Scene::GetTransformWorldMatrices() const {
	FlushMatrixSets();
	if (!worlds_valid) {
		if (ctx && !impl->topology_dirty)
			download_worlds(ctx, transform_worlds, transforms.capacity(), false);
		worlds_valid = true;
	}
	return transform_worlds;
}

const std::vector<Mat4> &Scene::GetPreviousTransformWorldMatrices() const {
	if (!previous_valid) {
		if (ctx && !impl->topology_dirty)
			download_worlds(ctx, previous_transform_worlds, transforms.capacity(), true);
		previous_valid = true;
	}
	return previous_transform_worlds;
}

// ---- bounding boxes ----------------------------------------------------------------------------------------------------------
// Scene::GetMinMax (scene.cpp:1982) = GetNodesMinMax(GetNodesWithComponent(NCI_Object)) (node.cpp:265-285): the per-node part
// (world * Union(model bounds), node.cpp:106-119 + component.cpp:778-797) and the union run on the device (hgb_minmax) over the
// scene's current world matrices; the host only lists which (transform slot, model) pairs take part, like the reference's node walk.
std::vector<Node> Scene::GetNodesWithObject() const {
	std::vector<Node> out;
	for (uint32_t ni = nodes.first(); ni != nodes.invalid_idx; ni = nodes.next(ni))
		if (objects.is_valid(nodes[ni].object))
			out.push_back({const_cast<Scene *>(this), nodes.get_ref(ni)});
	return out;
}

bool Scene::GetObjectMinMax(ComponentRef ref, const PipelineResources &resources, MinMax &minmax) const {
	if (!objects.is_valid(ref)) {
		warn("Invalid object component");
		return false;
	}
	const Object_ &obj = objects[ref.idx];
	if (!resources.models.IsValidRef(obj.model))
		return false;
	const Model &mdl = resources.models.Get_unsafe_(uint16_t(obj.model.ref.idx));
	if (mdl.bounds.empty())
		return false;
	minmax = mdl.bounds[0];
	for (size_t i = 1; i < mdl.bounds.size(); ++i) { // Union (minmax.h:48): Min/Max per component, `a < b ? a : b`
		const MinMax &b = mdl.bounds[i];
		minmax.mn = Vec3(minmax.mn.x < b.mn.x ? minmax.mn.x : b.mn.x, minmax.mn.y < b.mn.y ? minmax.mn.y : b.mn.y, minmax.mn.z < b.mn.z ? minmax.mn.z : b.mn.z);
		minmax.mx = Vec3(minmax.mx.x > b.mx.x ? minmax.mx.x : b.mx.x, minmax.mx.y > b.mx.y ? minmax.mx.y : b.mx.y, minmax.mx.z > b.mx.z ? minmax.mx.z : b.mx.z);
	}
	return true;
}

bool GetNodesMinMax(const std::vector<Node> &nodes_, const PipelineResources &resources, MinMax &minmax) {
	if (nodes_.empty())
		return false;
	Scene *scene = nullptr;
	std::vector<uint32_t> slot, ob;
	std::vector<float> obounds6;
	std::map<uint32_t, uint32_t> ob_of_model;
	for (const Node &n : nodes_) {
		if (!n.scene || !n.scene->nodes.is_valid(n.ref))
			continue;
		if (scene && n.scene != scene) {
			warn("GetNodesMinMax: nodes of different scenes");
			continue;
		}
		scene = n.scene;
		const Scene::Node_ &node = scene->nodes[n.ref.idx];
		if (!scene->objects.is_valid(node.object))
			continue; // HasObject() false and no instance view: Node::GetMinMax returns false
		MinMax mm;
		if (!scene->GetObjectMinMax(node.object, resources, mm))
			continue;
		const uint32_t mdl = scene->objects[node.object.idx].model.ref.idx;
		auto it = ob_of_model.find(mdl);
		if (it == ob_of_model.end()) {
			it = ob_of_model.emplace(mdl, uint32_t(obounds6.size() / 6)).first;
			const float v[6] = {mm.mn.x, mm.mn.y, mm.mn.z, mm.mx.x, mm.mx.y, mm.mx.z};
			obounds6.insert(obounds6.end(), v, v + 6);
		}
		slot.push_back(scene->transforms.is_valid(node.transform) ? node.transform.idx : HGB_ROOT); // invalid transform: identity world
		ob.push_back(it->second);
	}
	if (!scene || slot.empty())
		return false;
	scene->FlushMatrixSets();
	if (!scene->ctx || scene->impl->topology_dirty) {
		warn("GetNodesMinMax: the scene has no world matrices on the device yet (Update first)");
		return false;
	}
	float out6[6];
	uint32_t count = 0;
	if (hgb_set_objects(scene->ctx, uint32_t(slot.size()), slot.data(), ob.data(), uint32_t(obounds6.size() / 6), obounds6.data(), nullptr, 1) ||
		hgb_minmax(scene->ctx) || hgb_get_minmax(scene->ctx, out6, &count)) {
		warn(hgb_last_error());
		return false;
	}
	if (!count)
		return false;
	minmax = MinMax(Vec3(out6[0], out6[1], out6[2]), Vec3(out6[3], out6[4], out6[5]));
	return true;
}

bool Scene::GetMinMax(const PipelineResources &resources, MinMax &minmax) const { return GetNodesMinMax(GetNodesWithObject(), resources, minmax); }
bool Node::GetMinMax(const PipelineResources &resources, MinMax &minmax) const { return GetNodesMinMax({*this}, resources, minmax); }

// ---- display lists (host walk, like the reference: it only reads flags and refs, scene.cpp:318-404) -----------------------------
void Scene::GetModelDisplayLists(std::vector<ModelDisplayList> &out_opaque, std::vector<ModelDisplayList> &out_transparent,
	std::vector<SkinnedModelDisplayList> &out_opaque_skinned, std::vector<SkinnedModelDisplayList> &out_transparent_skinned,
	const PipelineResources &resources) const {
	RefreshFlagsFromDevice();
	out_opaque.clear(), out_transparent.clear(), out_opaque_skinned.clear(), out_transparent_skinned.clear();
	out_opaque.reserve(nodes.size());
	for (uint32_t ni = nodes.first(); ni != nodes.invalid_idx; ni = nodes.next(ni)) {
		const Node_ &node = nodes[ni];
		if (node.flags & (NF_Disabled | NF_InstanceDisabled))
			continue;
		if (!transforms.is_valid(node.transform) || !objects.is_valid(node.object))
			continue; // a valid transform is required even for skinned objects, like the reference (:335-336)
		const Object_ &obj = objects[node.object.idx];
		const uint16_t mdl_idx = resources.models.GetValidatedRefIndex(obj.model);
		const Model &mdl = resources.models.Get_unsafe_(mdl_idx);
		const size_t bone_count = obj.bones.size();
		const bool skinned = bone_count > 0 && bone_count == mdl.bind_pose.size();
		for (size_t i = 0; i < mdl.lists.size(); ++i) {
			const uint16_t mat_idx = mdl.mats[i];
			if (mat_idx >= obj.materials.size())
				continue; // silently dropped (:352)
			const Material *mat = &obj.materials[mat_idx];
			const bool transparent = GetMaterialBlendMode(*mat) != BM_Opaque;
			if (!skinned) {
				(transparent ? out_transparent : out_opaque).push_back({mat, node.transform.idx, mdl_idx, uint16_t(i)});
				continue;
			}
			SkinnedModelDisplayList dl{};
			dl.mat = mat;
			const std::vector<uint16_t> &table = mdl.lists[i].bones_table;
			const size_t n = std::min<size_t>(table.size(), max_skinned_model_matrix_count);
			for (size_t j = 0; j < n; ++j) {
				uint16_t bone_idx = table[j];
				uint32_t mtx_idx = node.transform.idx; // a stale bone ref falls back to the object's own transform (:372-383)
				if (bone_idx < bone_count) {
					const NodeRef bone = obj.bones[bone_idx];
					if (nodes.is_valid(bone) && transforms.is_valid(nodes[bone.idx].transform))
						mtx_idx = nodes[bone.idx].transform.idx;
				} else {
					bone_idx = 0;
				}
				dl.mtx_idxs[j] = mtx_idx;
				dl.bones_idxs[j] = bone_idx;
			}
			dl.bone_count = uint16_t(n);
			dl.mdl_idx = mdl_idx;
			dl.lst_idx = uint16_t(i);
			(transparent ? out_transparent_skinned : out_opaque_skinned).push_back(dl);
		}
	}
}

static Scene *scene_of(const std::vector<Mat4> &mtxs) {
	std::lock_guard<std::mutex> lock(g_registry_mutex);
	auto i = g_registry.find(&mtxs);
	return i == g_registry.end() ? nullptr : i->second;
}

// render_pipeline.cpp:2098-2104 through hgb_cull_list: any caller-owned list, any frustum, the scene's current world matrices
void CullModelDisplayLists(const Frustum &frustum, std::vector<ModelDisplayList> &display_lists, const std::vector<Mat4> &mtxs, const PipelineResources &resources) {
	if (display_lists.empty())
		return;
	Scene *scene = scene_of(mtxs);
	if (!scene || !scene->ctx) {
		warn("CullModelDisplayLists: the matrices do not belong to a live hg::Scene with a device context; list left unculled");
		return;
	}
	std::vector<float> bounds6;
	std::vector<uint32_t> base_of_model;
	build_bounds(resources, bounds6, base_of_model);
	const uint32_t n = uint32_t(display_lists.size());
	std::vector<uint32_t> mtx(n), bidx(n), out(n);
	for (uint32_t i = 0; i < n; ++i) {
		mtx[i] = display_lists[i].mtx_idx;
		bidx[i] = base_of_model[display_lists[i].mdl_idx] + display_lists[i].lst_idx;
	}
	uint32_t count = 0;
	if (hgb_cull_list(scene->ctx, n, mtx.data(), bidx.data(), uint32_t(bounds6.size() / 6), bounds6.data(), &frustum[0].x, out.data(), &count)) {
		warn(hgb_last_error());
		return;
	}
	for (uint32_t i = 0; i < count; ++i) // ascending survivor indices: the in-place gather never overwrites an unread entry
		display_lists[i] = display_lists[out[i]];
	display_lists.resize(count);
}

// render_pipeline.cpp:2157-2188 (math :2172-2173, previous frame :2177-2181) through hgb_set_skins / hgb_bone_matrices
size_t ComputeSkinMatrices(const std::vector<SkinnedModelDisplayList> &display_lists, const std::vector<Mat4> &mtxs, const PipelineResources &resources,
	std::vector<float> &out16, bool previous) {
	out16.clear();
	Scene *scene = scene_of(mtxs);
	if (!scene || !scene->ctx) {
		warn("ComputeSkinMatrices: the matrices do not belong to a live hg::Scene with a device context");
		return 0;
	}
	// bind-pose pool: every model's bind pose back to back
	std::vector<float> pool;
	std::vector<uint32_t> base_of_model(0x10000, 0);
	for (uint32_t m = resources.models.first(); m != 0xffffffffu; m = resources.models.next(m)) {
		const Model &mdl = resources.models.Get_unsafe_(uint16_t(m));
		base_of_model[m] = uint32_t(pool.size() / 12);
		for (const Mat4 &b : mdl.bind_pose)
			pool.insert(pool.end(), &b.m[0][0], &b.m[0][0] + 12);
	}
	std::vector<uint32_t> mtx, bind;
	for (const SkinnedModelDisplayList &dl : display_lists)
		for (uint16_t j = 0; j < dl.bone_count; ++j) {
			mtx.push_back(dl.mtx_idxs[j]);
			bind.push_back(base_of_model[dl.mdl_idx] + dl.bones_idxs[j]);
		}
	if (mtx.empty())
		return 0;
	Scene::Impl *impl = scene->impl;
	if (mtx != impl->skin_mtx || bind != impl->skin_bind || pool.size() != impl->skin_bind_count) { // records changed: re-upload
		if (hgb_set_skins(scene->ctx, uint32_t(mtx.size()), mtx.data(), bind.data(), uint32_t(pool.size() / 12), pool.data())) {
			warn(hgb_last_error());
			return 0;
		}
		impl->skin_mtx = mtx, impl->skin_bind = bind, impl->skin_bind_count = pool.size();
	}
	out16.resize(mtx.size() * 16);
	if (hgb_bone_matrices(scene->ctx, previous ? 1 : 0) || hgb_get_bone_matrices(scene->ctx, out16.data(), previous ? 1 : 0)) {
		warn(hgb_last_error());
		out16.clear();
		return 0;
	}
	return mtx.size();
}

// opt-in cull of skinned lists (see render_pipeline.h): hgb_cull_skinned_list over the caller's vector
void CullSkinnedModelDisplayLists(const Frustum &frustum, std::vector<SkinnedModelDisplayList> &display_lists, const std::vector<Mat4> &mtxs,
	const PipelineResources &resources) {
	if (display_lists.empty())
		return;
	Scene *scene = scene_of(mtxs);
	if (!scene || !scene->ctx) {
		warn("CullSkinnedModelDisplayLists: the matrices do not belong to a live hg::Scene with a device context; list left unculled");
		return;
	}
	std::vector<float> pool, bounds6;
	std::vector<uint32_t> bind_base(0x10000, 0), bounds_base;
	build_bounds(resources, bounds6, bounds_base);
	for (uint32_t m = resources.models.first(); m != 0xffffffffu; m = resources.models.next(m)) {
		const Model &mdl = resources.models.Get_unsafe_(uint16_t(m));
		bind_base[m] = uint32_t(pool.size() / 12);
		for (const Mat4 &b : mdl.bind_pose)
			pool.insert(pool.end(), &b.m[0][0], &b.m[0][0] + 12);
	}
	std::vector<uint32_t> off{0}, mtx, bind, bidx;
	for (const SkinnedModelDisplayList &dl : display_lists) {
		for (uint16_t j = 0; j < dl.bone_count; ++j) {
			mtx.push_back(dl.mtx_idxs[j]);
			bind.push_back(bind_base[dl.mdl_idx] + dl.bones_idxs[j]);
		}
		off.push_back(uint32_t(mtx.size()));
		bidx.push_back(bounds_base[dl.mdl_idx] + dl.lst_idx);
	}
	if (pool.empty() || bounds6.empty())
		return;
	std::vector<uint32_t> keep(display_lists.size());
	uint32_t n = 0;
	if (hgb_cull_skinned_list(scene->ctx, uint32_t(display_lists.size()), off.data(), mtx.data(), bind.data(), uint32_t(pool.size() / 12), pool.data(), bidx.data(),
			uint32_t(bounds6.size() / 6), bounds6.data(), &frustum[0].x, keep.data(), &n)) {
		warn(hgb_last_error());
		return;
	}
	for (uint32_t i = 0; i < n; ++i) // ascending indices: an in-place stable compaction, like std::remove_if
		if (keep[i] != i)
			display_lists[i] = display_lists[keep[i]];
	display_lists.resize(n);
}

// ---- garbage collection (scene.cpp:92-237): components no node refers to, anims no scene anim refers to -----------------------
size_t Scene::GarbageCollect() {
	size_t removed = 0;
	std::vector<uint8_t> trs_used(transforms.capacity(), 0), obj_used(objects.capacity(), 0);
	for (uint32_t ni = nodes.first(); ni != nodes.invalid_idx; ni = nodes.next(ni)) {
		if (transforms.is_valid(nodes[ni].transform))
			trs_used[nodes[ni].transform.idx] = 1;
		if (objects.is_valid(nodes[ni].object))
			obj_used[nodes[ni].object.idx] = 1;
	}
	for (uint32_t i = transforms.first(); i != transforms.invalid_idx;) {
		const uint32_t nx = transforms.next(i);
		if (!trs_used[i]) {
			transforms.remove_ref(transforms.get_ref(i));
			++removed;
			impl->topology_dirty = true, impl->topology_reparent_only = impl->topology_append_only = false;
		}
		i = nx;
	}
	for (uint32_t i = objects.first(); i != objects.invalid_idx;) {
		const uint32_t nx = objects.next(i);
		if (!obj_used[i]) {
			objects.remove_ref(objects.get_ref(i));
			++removed;
			impl->touch_entries_all();
		}
		i = nx;
	}
	std::vector<uint8_t> cam_used(cameras.capacity(), 0); // scene.cpp:92-237 runs the same pass over every component list
	for (uint32_t ni = nodes.first(); ni != nodes.invalid_idx; ni = nodes.next(ni))
		if (cameras.is_valid(nodes[ni].camera))
			cam_used[nodes[ni].camera.idx] = 1;
	for (uint32_t i = cameras.first(); i != cameras.invalid_idx;) {
		const uint32_t nx = cameras.next(i);
		if (!cam_used[i]) {
			cameras.remove_ref(cameras.get_ref(i));
			++removed;
		}
		i = nx;
	}
	std::vector<uint8_t> lgt_used(lights.capacity(), 0);
	for (uint32_t ni = nodes.first(); ni != nodes.invalid_idx; ni = nodes.next(ni))
		if (lights.is_valid(nodes[ni].light))
			lgt_used[nodes[ni].light.idx] = 1;
	for (uint32_t i = lights.first(); i != lights.invalid_idx;) {
		const uint32_t nx = lights.next(i);
		if (!lgt_used[i]) {
			lights.remove_ref(lights.get_ref(i));
			++removed;
		}
		i = nx;
	}
	std::vector<uint8_t> inst_used(instances.capacity(), 0);
	for (const auto &ni : node_instance)
		if (nodes.is_valid(ni.first) && instances.is_valid(ni.second))
			inst_used[ni.second.idx] = 1;
	for (uint32_t i = instances.first(); i != instances.invalid_idx;) {
		const uint32_t nx = instances.next(i);
		if (!inst_used[i]) {
			instances.remove_ref(instances.get_ref(i));
			++removed;
		}
		i = nx;
	}
	if (removed)
		impl->anims_dirty = true;
	return removed;
}

} // namespace hg
