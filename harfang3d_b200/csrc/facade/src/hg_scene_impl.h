// internal: dirty-state / staging block of hg::Scene (shared by the facade's translation units)
#pragma once
#include "engine/scene.h"
#include <map>
#include <string>
#include <vector>

namespace hg {

struct Scene::Impl {
	bool topology_dirty{true};
	bool topology_reparent_only{false}; // since the last sync the only hierarchy edits were SetTransformParent calls on `reparented`
	std::vector<uint32_t> reparented;
	// In-place spawning (hgb_options.append_slack, Scene::SetDefaultAppendSlack): while the only hierarchy edits since the last sync are
	// CreateTransform calls and SetTransformParent / SetNodeTransform on those fresh transforms, the sync appends them (hgb_append_nodes)
	// instead of re-sorting the hierarchy; while the only display-list edits touch nodes behind the last node of the entry tables, it
	// appends their entries; Enable / DisableNode flips entry flags in place.
	int append_slack{0};                     // Scene::SetDefaultAppendSlack at construction
	bool topology_append_only{false};
	bool entries_append_only{false};
	uint32_t ent_nodes_end{0};               // node slots [0, ent_nodes_end) were walked by the last entry-table build
	bool ent_built{false};
	size_t ent_bounds_count{0};              // bounds rows the device holds
	std::vector<uint32_t> ent_node[2];       // node index of every entry (ascending): a node's entries are found by binary search
	std::vector<uint32_t> flag_nodes;        // nodes whose enabled state may have flipped since the last sync
	// "created since the last sync": an epoch stamp per transform slot / node slot (no per-sync clearing, O(1) lookup)
	uint32_t epoch{1};
	std::vector<uint32_t> slot_epoch, node_epoch;
	static void stamp(std::vector<uint32_t> &v, uint32_t idx, uint32_t e) {
		if (idx >= v.size())
			v.resize(size_t(idx) + 1 + v.size() / 2, 0u);
		v[idx] = e;
	}
	bool fresh_slot(uint32_t idx) const { return idx < slot_epoch.size() && slot_epoch[idx] == epoch; }
	bool fresh_node(uint32_t idx) const { return idx < node_epoch.size() && node_epoch[idx] == epoch; }
	void touch_entries(uint32_t node_idx) {  // an edit that changes the display entries of ONE node
		if (!entries_dirty)
			entries_append_only = true;
		if (node_idx < ent_nodes_end)
			entries_append_only = false;
		entries_dirty = true;
	}
	void touch_entries_all() { entries_dirty = true, entries_append_only = false; }
	bool trs_all_dirty{true};
	std::vector<uint32_t> dirty_trs;
	bool anims_dirty{true};
	// Single-value animation targets are evaluated on the host every frame, in the reference's evaluation order (EvaluateBoundAnim,
	// scene.cpp:2183-2215): one float / colour / vec4 per component — no kernel. `clock` says whose time drives the track: a play slot.
	enum HostTarget : uint8_t { HT_CameraFov, HT_LightDiffuse, HT_LightSpecular, HT_LightDiffuseIntensity, HT_LightSpecularIntensity, HT_MaterialValue, HT_EnableInstanceHost };
	struct HostTrack {
		HostTarget target;
		uint32_t clock;     // play slot whose clock (impl->play_t) drives it
		ComponentRef comp;  // camera / light / object (material) ; HT_EnableInstanceHost: unused
		NodeRef node;
		AnimRef anim;
		int track;
		uint8_t mat_slot{0};
		std::string mat_value;
	};
	std::vector<HostTrack> host_tracks; // evaluation order, rebuilt with the bindings
	// Nested instance anims (Anim::instance_anim_track, scene.cpp:2217-2268): the node anim of a playing (or itself nested) scene anim
	// selects, by a step track, which scene anim of the node's instantiated scene is evaluated, at a clock derived from the parent's.
	// Every such evaluation owns a play slot of the device; the host derives its clock each frame (exact integer arithmetic of the
	// reference) and re-binds when the step track moves to another key.
	struct Nested {
		int parent;          // index into `nested`, -1 = driven by a playing anim
		uint32_t play_idx;   // play_anims index when parent == -1
		NodeRef node;
		AnimRef anim;        // the anim holding the instance_anim_track
		int kf{-2};          // key in force when the tracks were bound (-1 = before the first key)
		SceneAnimRef sub;    // scene anim of the instance view in force (invalid = nothing evaluated)
		uint32_t play_slot{0};
	};
	std::vector<Nested> nested;
	uint32_t n_play_slots{0};
	bool entries_dirty{true};
	bool views_dirty{false};
	bool in_frame{false}; // between ReadyWorldMatrices and ComputeWorldMatrices: SetNodeWorldMatrix records presets
	const PipelineResources *resources{nullptr};
	std::vector<Frustum> frusta;
	std::vector<uint32_t> preset_slots;
	std::vector<float> preset_mats;
	std::vector<std::pair<uint32_t, uint32_t>> enable_nodes; // (transform slot, node index) of the nodes driven by an "Enable" track
	bool enable_state_push{false};     // the host's flags of the tracked nodes must be pushed (tracks were (re)bound or entries rebuilt)
	bool flags_stale{false};           // the device evaluated Enable tracks: NF_Disabled of those nodes is refreshed before the host reads it
	uint32_t device_slots{0};          // slots of the topology the device currently holds
	std::vector<uint32_t> fresh_slots; // transforms created since then: their host TRS is newer than whatever the device has in that slot
	std::vector<uint32_t> set_slots; // queued SetTransformWorld/LocalMatrix calls (one mode per queue)
	std::vector<float> set_mats;
	bool set_local{false};
	std::vector<int64_t> play_t;        // evaluation time per play slot (HGB_NOT_PLAYING when idle)
	std::vector<float> staging;
	// fused culling: the entry tables given to the device (every entry, disabled ones flagged) and what they map back to
	std::vector<ModelDisplayList> ent[2];
	std::vector<uint8_t> ent_flags[2];
	// ComputeSkinMatrices cache
	std::vector<uint32_t> skin_mtx, skin_bind;
	size_t skin_bind_count{0};
	bool warned_no_device{false};
	bool rebind_after_frame{false}; // a finished ALM_Once anim is evaluated one last time, then its tracks are unbound
};


} // namespace hg
