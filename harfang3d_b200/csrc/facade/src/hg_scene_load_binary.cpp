// hg::Scene facade — scene binary (format version 9) and its embedded animation chunk, read straight into the facade's containers in
// bulk: no per-node Scene API calls, no SetParent cycle walks; the next frame uploads the whole hierarchy as one SoA batch
// (hgb_set_topology + hgb_upload_trs + entry / track tables). SURVEY.md §8f-3.
//
// Byte layout follows what the reference WRITES (harfang/engine/scene_load_binary.cpp:237-541 Save_binary, :16-98 SaveComponent,
// harfang/engine/render_pipeline.cpp:1317-1349 SaveMaterial, harfang/engine/anim_load_binary.cpp:42-98 SaveAnim*, strings =
// u16 length + bytes, harfang/foundation/rw_interface.cpp:34-37) and is consumed in the order Load_binary reads it (:549-905).
// Components this path has no use for (physics, scripts, environment, key/values) are parsed and dropped; lights and instances are kept
// (animation targets, NF_InstanceDisabled propagation, nested instance anims).
// Resources are never loaded from disk here: a model name resolves to the model registered under that name in `resources`, or to an
// empty placeholder registered now (LSSF_DoNotLoadResources behaviour, render_pipeline.cpp:1669-1672).
#include "engine/scene.h"
#include "foundation/log.h"

#include <algorithm>
#include <cstdio>
#include <cstring>

namespace hg {

namespace {

struct Cursor {
	const uint8_t *p, *end;
	bool ok{true};
	template <typename T> T get() {
		T v{};
		if (size_t(end - p) < sizeof(T)) {
			ok = false, p = end;
			return v;
		}
		memcpy(&v, p, sizeof(T));
		p += sizeof(T);
		return v;
	}
	void skip(size_t n) {
		if (size_t(end - p) < n)
			ok = false, p = end;
		else
			p += n;
	}
	std::string str() {
		const uint16_t n = get<uint16_t>();
		if (size_t(end - p) < n) {
			ok = false, p = end;
			return {};
		}
		std::string s(reinterpret_cast<const char *>(p), n);
		p += n;
		const size_t z = s.find('\0'); // the reference reads into a C string (rw_interface.cpp:23-28)
		if (z != std::string::npos)
			s.resize(z);
		return s;
	}
	Vec3 vec3() {
		const float x = get<float>(), y = get<float>(), z = get<float>();
		return Vec3(x, y, z);
	}
};

// bgfx render state -> blend mode (render_pipeline.cpp:394-417): four 4-bit blend factors at bit 12 (src/dst RGB, src/dst alpha), two 3-bit
// equations at bit 28. Factor codes: 1 zero, 2 one, 4 1-src colour, 5 src alpha, 6 1-src alpha, 9 dst colour, 10 1-dst colour;
// equations: 0 add, 1 sub, 3 min, 4 max.
BlendMode blend_mode_of_state(uint64_t state) {
	const uint32_t factors = uint32_t(state >> 12) & 0xffffu, eq = uint32_t(state >> 28) & 0x3fu;
	if (factors == 0x2222u && eq == 0) return BM_Additive;
	if (factors == 0x6565u && eq == 0) return BM_Alpha;
	if (factors == 0x2222u && eq == (3u | 3u << 3)) return BM_Darken;
	if (factors == 0x2222u && eq == (4u | 4u << 3)) return BM_Lighten;
	if (factors == 0x1919u && eq == 0) return BM_Multiply;
	if (factors == 0x6262u && eq == 0) return BM_Opaque;
	if (factors == 0x4242u && eq == 0) return BM_Screen;
	if (factors == 0xa9a9u && eq == (1u | 1u << 3)) return BM_LinearBurn;
	if (factors == 0x2265u && eq == 0) return BM_AlphaRGB_AddAlpha;
	return BM_Opaque;
}

Material read_material(Cursor &c) {
	c.str(); // program name
	const uint16_t value_count = c.get<uint16_t>();
	for (uint16_t i = 0; i < value_count && c.ok; ++i) {
		c.str();
		c.skip(4 + 2); // uniform type (enum, 4 B), count
		const uint16_t n = c.get<uint16_t>();
		c.skip(size_t(n) * 4);
	}
	const uint16_t texture_count = c.get<uint16_t>();
	for (uint16_t i = 0; i < texture_count && c.ok; ++i) {
		c.str();
		c.skip(1); // channel
		if (c.get<uint8_t>() == 1) {
			c.str();
			c.skip(4);
		}
	}
	const uint64_t state = c.get<uint64_t>();
	c.skip(4 + 1); // rgba, flags
	Material m;
	m.blend_mode = blend_mode_of_state(state);
	return m;
}

template <typename Track> void sort_keys(Track &track) { // SortAnimTrackKeys, animation.h:172-177
	std::stable_sort(track.keys.begin(), track.keys.end(), [](const typename decltype(track.keys)::value_type &a, const typename decltype(track.keys)::value_type &b) { return a.t < b.t; });
	track.keys.erase(std::unique(track.keys.begin(), track.keys.end(), [](const typename decltype(track.keys)::value_type &a, const typename decltype(track.keys)::value_type &b) { return a.t == b.t; }),
		track.keys.end());
}

// tracks of a type the facade does not evaluate: target + count x (time + payload [+ tension, bias])
void skip_tracks(Cursor &c, size_t key_bytes) {
	const uint32_t n = c.get<uint32_t>();
	for (uint32_t i = 0; i < n && c.ok; ++i) {
		c.str();
		const uint32_t k = c.get<uint32_t>();
		c.skip(size_t(k) * key_bytes);
	}
}

void read_anim(Cursor &c, Anim &anim) { // LoadAnimFromBinary, anim_load_binary.cpp:181-208
	const uint16_t version = c.get<uint16_t>();
	if (version > 2) {
		warn("Unsupported animation format version");
		c.ok = false;
		return;
	}
	anim.t_start = c.get<int64_t>();
	anim.t_end = c.get<int64_t>();
	anim.flags = c.get<uint8_t>();
	{ // bool tracks ("Enable")
		const uint32_t n = c.get<uint32_t>();
		anim.bool_tracks.resize(c.ok ? n : 0);
		for (auto &t : anim.bool_tracks) {
			t.target = c.str();
			const uint32_t k = c.get<uint32_t>();
			if (!c.ok || size_t(c.end - c.p) < size_t(k) * 9) {
				c.ok = false;
				return;
			}
			t.keys.resize(k);
			for (auto &key : t.keys)
				key.t = c.get<int64_t>(), key.v = c.get<uint8_t>() != 0;
			sort_keys(t);
		}
	}
	skip_tracks(c, 8 + 4);         // int
	{                              // float, Hermite: "Camera.Fov" (the light intensities have no component here)
		const uint32_t n = c.get<uint32_t>();
		anim.float_tracks.resize(c.ok ? n : 0);
		for (auto &t : anim.float_tracks) {
			t.target = c.str();
			const uint32_t k = c.get<uint32_t>();
			if (!c.ok || size_t(c.end - c.p) < size_t(k) * 20) {
				c.ok = false;
				return;
			}
			t.keys.resize(k);
			for (auto &key : t.keys) {
				key.t = c.get<int64_t>();
				key.v = c.get<float>();
				key.tension = c.get<float>(), key.bias = c.get<float>();
			}
			sort_keys(t);
		}
	}
	skip_tracks(c, 8 + 8 + 8);     // vec2, Hermite
	{                              // vec3, Hermite: "Position", "Rotation", "Scale"
		const uint32_t n = c.get<uint32_t>();
		anim.vec3_tracks.resize(c.ok ? n : 0);
		for (auto &t : anim.vec3_tracks) {
			t.target = c.str();
			const uint32_t k = c.get<uint32_t>();
			if (!c.ok || size_t(c.end - c.p) < size_t(k) * 28) {
				c.ok = false;
				return;
			}
			t.keys.resize(k);
			for (auto &key : t.keys) {
				key.t = c.get<int64_t>();
				key.v = c.vec3();
				key.tension = c.get<float>(), key.bias = c.get<float>();
			}
			sort_keys(t);
		}
	}
	{                           // vec4, Hermite: "Material.<slot>.<value>"
		const uint32_t n = c.get<uint32_t>();
		anim.vec4_tracks.resize(c.ok ? n : 0);
		for (auto &t : anim.vec4_tracks) {
			t.target = c.str();
			const uint32_t k = c.get<uint32_t>();
			if (!c.ok || size_t(c.end - c.p) < size_t(k) * 32) {
				c.ok = false;
				return;
			}
			t.keys.resize(k);
			for (auto &key : t.keys) {
				key.t = c.get<int64_t>();
				key.v.x = c.get<float>(), key.v.y = c.get<float>(), key.v.z = c.get<float>(), key.v.w = c.get<float>();
				key.tension = c.get<float>(), key.bias = c.get<float>();
			}
			sort_keys(t);
		}
	}
	{                           // quaternion, linear
		const uint32_t n = c.get<uint32_t>();
		anim.quat_tracks.resize(c.ok ? n : 0);
		for (auto &t : anim.quat_tracks) {
			t.target = c.str();
			const uint32_t k = c.get<uint32_t>();
			if (!c.ok || size_t(c.end - c.p) < size_t(k) * 24) {
				c.ok = false;
				return;
			}
			t.keys.resize(k);
			for (auto &key : t.keys) {
				key.t = c.get<int64_t>();
				key.v.x = c.get<float>(), key.v.y = c.get<float>(), key.v.z = c.get<float>(), key.v.w = c.get<float>();
			}
			sort_keys(t);
		}
	}
	{                           // colour, Hermite: "Light.Diffuse", "Light.Specular"
		const uint32_t n = c.get<uint32_t>();
		anim.color_tracks.resize(c.ok ? n : 0);
		for (auto &t : anim.color_tracks) {
			t.target = c.str();
			const uint32_t k = c.get<uint32_t>();
			if (!c.ok || size_t(c.end - c.p) < size_t(k) * 32) {
				c.ok = false;
				return;
			}
			t.keys.resize(k);
			for (auto &key : t.keys) {
				key.t = c.get<int64_t>();
				key.v.r = c.get<float>(), key.v.g = c.get<float>(), key.v.b = c.get<float>(), key.v.a = c.get<float>();
				key.tension = c.get<float>(), key.bias = c.get<float>();
			}
			sort_keys(t);
		}
	}
	{                           // string tracks: variable-size keys
		const uint32_t n = c.get<uint32_t>();
		for (uint32_t i = 0; i < n && c.ok; ++i) {
			c.str();
			const uint32_t k = c.get<uint32_t>();
			for (uint32_t j = 0; j < k && c.ok; ++j) {
				c.skip(8);
				c.str();
			}
		}
	}
	if (version >= 2) { // instance anim track
		const uint32_t k = c.get<uint32_t>();
		for (uint32_t j = 0; j < k && c.ok; ++j) {
			AnimKeyT<InstanceAnimKey> key;
			key.t = c.get<int64_t>();
			key.v.anim_name = c.str();
			key.v.loop_mode = AnimLoopMode(c.get<uint8_t>());
			key.v.t_scale = c.get<float>();
			if (c.ok)
				anim.instance_anim_track.keys.push_back(std::move(key));
		}
		sort_keys(anim.instance_anim_track);
	}
}

} // namespace

uint32_t GetSceneBinaryFormatVersion() { return 9; }

bool Scene::Load_binary(const void *data, size_t size, const char *name, PipelineResources &resources, const PipelineInfo &pipeline, LoadSceneContext &ctx, uint32_t load_flags) {
	Cursor c{static_cast<const uint8_t *>(data), static_cast<const uint8_t *>(data) + size};
	char msg[256];
	auto refuse = [&](const char *why) {
		snprintf(msg, sizeof(msg), "Cannot load scene '%s', %s", name ? name : "", why);
		warn(msg);
		return false;
	};
	if (!data || !size)
		return refuse("invalid read handle");
	if (c.get<uint32_t>() != 0x46464748u) // 'HGFF'
		return refuse("invalid magic marker");
	if (c.get<uint8_t>() != 0x01)
		return refuse("invalid scene marker");
	if (c.get<uint32_t>() != GetSceneBinaryFormatVersion())
		return refuse("unsupported binary version");
	const uint32_t file_flags = c.get<uint32_t>();

	// ---- components ----
	const uint32_t transform_count = c.get<uint32_t>();
	if (!c.ok || size_t(c.end - c.p) < size_t(transform_count) * 40)
		return refuse("truncated transform table");
	std::vector<ComponentRef> transform_refs(transform_count);
	transforms.reserve(transforms.capacity() + transform_count);
	for (uint32_t i = 0; i < transform_count; ++i) {
		Transform_ t;
		t.TRS.pos = c.vec3(), t.TRS.rot = c.vec3(), t.TRS.scl = c.vec3();
		t.parent = NodeRef{c.get<uint32_t>(), 0xffffffffu}; // saved node index; remapped below. gen is irrelevant until then
		transform_refs[i] = transforms.add_ref(t);
	}
	const uint32_t camera_count = c.get<uint32_t>();
	if (!c.ok || size_t(c.end - c.p) < size_t(camera_count) * 17)
		return refuse("truncated camera table");
	std::vector<ComponentRef> camera_refs(camera_count);
	for (uint32_t i = 0; i < camera_count; ++i) { // zrange 8, fov 4, ortho 1, size 4 (scene_load_binary.cpp: SaveComponent(Camera_))
		Camera_ cam;
		cam.zrange.znear = c.get<float>(), cam.zrange.zfar = c.get<float>();
		cam.fov = c.get<float>();
		cam.ortho = c.get<uint8_t>() != 0;
		cam.size = c.get<float>();
		camera_refs[i] = cameras.add_ref(cam);
	}
	const uint32_t object_count = c.get<uint32_t>();
	std::vector<ComponentRef> object_refs;
	object_refs.reserve(std::min<size_t>(object_count, size_t(c.end - c.p)));
	for (uint32_t i = 0; i < object_count && c.ok; ++i) {
		Object_ o;
		const std::string model_name = c.str();
		if (!model_name.empty()) {
			o.model = resources.models.Has(model_name);
			if (!resources.models.IsValidRef(o.model))
				o.model = resources.models.Add(model_name, Model{}); // placeholder: geometry files are not read on this path
		}
		const uint16_t mat_count = c.get<uint16_t>();
		o.materials.resize(c.ok ? mat_count : 0);
		for (auto &m : o.materials)
			m = read_material(c);
		for (uint16_t j = 0; j < mat_count && c.ok; ++j)
			c.str(); // material_infos[j].name
		const uint16_t bone_count = c.get<uint16_t>();
		if (size_t(c.end - c.p) < size_t(bone_count) * 4)
			c.ok = false;
		o.bones.resize(c.ok ? bone_count : 0);
		for (auto &b : o.bones)
			b = NodeRef{c.get<uint32_t>(), 0xffffffffu};
		object_refs.push_back(objects.add_ref(std::move(o)));
	}
	const uint32_t light_count = c.get<uint32_t>(); // scene_load_binary.cpp:148-161: 84 bytes each
	if (size_t(c.end - c.p) < size_t(light_count) * 84)
		c.ok = false;
	std::vector<ComponentRef> light_refs(c.ok ? light_count : 0);
	for (uint32_t i = 0; i < light_count && c.ok; ++i) {
		Light_ l;
		l.type = LightType(c.get<uint32_t>());
		l.shadow_type = LightShadowType(c.get<uint32_t>());
		l.diffuse.r = c.get<float>(), l.diffuse.g = c.get<float>(), l.diffuse.b = c.get<float>(), l.diffuse.a = c.get<float>();
		l.diffuse_intensity = c.get<float>();
		l.specular.r = c.get<float>(), l.specular.g = c.get<float>(), l.specular.b = c.get<float>(), l.specular.a = c.get<float>();
		l.specular_intensity = c.get<float>();
		l.radius = c.get<float>(), l.inner_angle = c.get<float>(), l.outer_angle = c.get<float>();
		l.pssm_split.x = c.get<float>(), l.pssm_split.y = c.get<float>(), l.pssm_split.z = c.get<float>(), l.pssm_split.w = c.get<float>();
		l.priority = c.get<float>(), l.shadow_bias = c.get<float>();
		light_refs[i] = lights.add_ref(l);
	}
	if (file_flags & LSSF_Physics) {
		const uint32_t rigid_body_count = c.get<uint32_t>();
		c.skip(size_t(rigid_body_count) * 6);
		const uint32_t collision_count = c.get<uint32_t>();
		for (uint32_t i = 0; i < collision_count && c.ok; ++i) {
			c.skip(1 + 4);
			c.str();
			c.skip(36);
		}
	}
	if (file_flags & LSSF_Scripts) {
		const uint32_t script_count = c.get<uint32_t>();
		for (uint32_t i = 0; i < script_count && c.ok; ++i) {
			c.str();
			const uint16_t n = c.get<uint16_t>();
			for (uint16_t j = 0; j < n && c.ok; ++j) {
				c.str();
				const uint16_t type = c.get<uint16_t>(); // SPT_Null, SPT_Int, SPT_Float, SPT_String, SPT_Bool
				if (type == 1 || type == 2)
					c.skip(4);
				else if (type == 3)
					c.str();
				else if (type == 4)
					c.skip(1);
			}
		}
	}
	const uint32_t instance_count = c.get<uint32_t>(); // scene_load_binary.cpp:203-207
	std::vector<ComponentRef> instance_refs;
	for (uint32_t i = 0; i < instance_count && c.ok; ++i) {
		Instance_ inst;
		inst.name = c.str();
		inst.anim = c.str();
		inst.loop_mode = AnimLoopMode(c.get<uint8_t>());
		instance_refs.push_back(instances.add_ref(std::move(inst)));
	}
	if (!c.ok)
		return refuse("truncated component tables");

	// ---- nodes ----
	std::vector<NodeRef> nodes_to_disable, nodes_with_instance;
	if (file_flags & LSSF_Nodes) {
		const uint32_t node_count = c.get<uint32_t>();
		nodes.reserve(nodes.capacity() + std::min<size_t>(node_count, size_t(c.end - c.p)));
		for (uint32_t i = 0; i < node_count && c.ok; ++i) {
			const uint32_t saved_idx = c.get<uint32_t>();
			Node_ n;
			n.name = c.str();
			const uint32_t node_flags = c.get<uint32_t>();
			const uint32_t transform_idx = c.get<uint32_t>();
			const uint32_t camera_idx = c.get<uint32_t>();
			const uint32_t object_idx = c.get<uint32_t>();
			const uint32_t light_idx = c.get<uint32_t>();
			if (light_idx != 0xffffffffu && light_idx < light_refs.size())
				n.light = light_refs[light_idx];
			if (transform_idx != 0xffffffffu && transform_idx < transform_refs.size())
				n.transform = transform_refs[transform_idx];
			if (object_idx != 0xffffffffu && object_idx < object_refs.size())
				n.object = object_refs[object_idx];
			if (camera_idx != 0xffffffffu && camera_idx < camera_refs.size())
				n.camera = camera_refs[camera_idx];
			if (file_flags & LSSF_Physics) {
				c.skip(4); // rigid body
				const uint32_t n_col = c.get<uint32_t>();
				c.skip(size_t(n_col) * 4);
			}
			if (file_flags & LSSF_Scripts) {
				const uint32_t n_scr = c.get<uint32_t>();
				c.skip(size_t(n_scr) * 4);
			}
			const uint32_t instance_idx = c.get<uint32_t>();
			if (!c.ok)
				break;
			const NodeRef ref = nodes.add_ref(std::move(n));
			if (instance_idx != 0xffffffffu && instance_idx < instance_refs.size()) {
				node_instance[ref] = instance_refs[instance_idx];
				nodes_with_instance.push_back(ref);
			}
			ctx.node_refs[saved_idx] = ref;
			ctx.view.nodes.push_back(ref);
			if (node_flags & NF_Disabled)
				nodes_to_disable.push_back(ref);
		}
		// fix parent and bone references: saved node index -> the node created for it (stale ones become invalid refs = roots)
		for (const ComponentRef &r : transform_refs) {
			Transform_ &t = transforms[r.idx];
			if (t.parent.idx != 0xffffffffu) {
				const auto i = ctx.node_refs.find(t.parent.idx);
				t.parent = i != ctx.node_refs.end() ? i->second : InvalidNodeRef;
			} else {
				t.parent = InvalidNodeRef;
			}
		}
		for (const ComponentRef &r : object_refs)
			for (NodeRef &b : objects[r.idx].bones) {
				if (b.idx != 0xffffffffu) {
					const auto i = ctx.node_refs.find(b.idx);
					b = i != ctx.node_refs.end() ? i->second : InvalidNodeRef;
				} else {
					b = InvalidNodeRef;
				}
			}
	} else {
		for (const ComponentRef &r : transform_refs)
			transforms[r.idx].parent = InvalidNodeRef;
	}

	// ---- instances (scene_load_binary.cpp:714-722): each instance node loads its scene file under itself, then starts its
	// on-instantiate anim. Like the reference, not under LSSF_DoNotLoadResources. ----
	if (!(load_flags & LSSF_DoNotLoadResources) && c.ok) {
		for (const NodeRef &ref : nodes_to_disable) // the host's state decides NF_InstanceDisabled of what it instantiates
			if (nodes.is_valid(ref))
				nodes[ref.idx].flags |= NF_Disabled;
		for (const NodeRef &ref : nodes_with_instance) {
			NodeSetupInstanceFromFile(ref, resources, pipeline, LSSF_All | (load_flags & LSSF_DoNotLoadResources), ctx.recursion_level + 1);
			NodeStartOnInstantiateAnim(ref);
		}
	}

	// ---- scene block (environment, canvas): nothing the update path reads ----
	if (file_flags & LSSF_Scene) {
		if (file_flags & LSSF_Nodes) { // current camera: a saved node index (scene_load_binary.cpp:752-757)
			const uint32_t camera_idx = c.get<uint32_t>();
			const auto it = ctx.node_refs.find(camera_idx);
			if (camera_idx != 0xffffffffu && it != ctx.node_refs.end())
				current_camera = it->second;
		}
		c.skip(16 + 4 + 4 + 16); // ambient, fog near / far, fog colour
		c.str(), c.str();        // probe maps
		c.skip(1 + 1 + 36);      // probe type, parallax, TRS
		c.str();                 // brdf map
		c.skip(1 + 1 + 16);      // clear z, clear colour, colour
	}

	// ---- animations ----
	if ((file_flags & LSSF_Anims) && c.ok) {
		const uint32_t chunk_size = c.get<uint32_t>();
		if (load_flags & LSSF_Anims) {
			std::map<uint32_t, AnimRef> anim_refs;
			const uint32_t anim_count = c.get<uint32_t>();
			for (uint32_t i = 0; i < anim_count && c.ok; ++i) {
				const uint32_t idx = c.get<uint32_t>();
				Anim anim;
				read_anim(c, anim);
				if (!c.ok)
					break;
				const AnimRef ref = anims.add_ref(std::move(anim));
				ctx.view.anims.push_back(ref);
				anim_refs[idx] = ref;
			}
			const uint32_t scene_anim_count = c.get<uint32_t>();
			for (uint32_t i = 0; i < scene_anim_count && c.ok; ++i) {
				SceneAnim sa;
				sa.name = c.str();
				sa.t_start = c.get<int64_t>();
				sa.t_end = c.get<int64_t>();
				const uint32_t scene_anim_idx = c.get<uint32_t>();
				sa.frame_duration = c.get<int64_t>();
				const auto isa = anim_refs.find(scene_anim_idx);
				if (isa != anim_refs.end())
					sa.scene_anim = isa->second;
				const uint32_t node_anim_count = c.get<uint32_t>();
				if (size_t(c.end - c.p) < size_t(node_anim_count) * 8)
					c.ok = false;
				for (uint32_t j = 0; j < node_anim_count && c.ok; ++j) {
					const uint32_t node_idx = c.get<uint32_t>(), anim_idx = c.get<uint32_t>();
					const auto in = ctx.node_refs.find(node_idx);
					const auto ia = anim_refs.find(anim_idx);
					if (in != ctx.node_refs.end() && ia != anim_refs.end())
						sa.node_anims.push_back({in->second, ia->second});
				}
				if (!c.ok)
					break;
				ctx.view.scene_anims.push_back(scene_anims.add_ref(std::move(sa)));
			}
		} else {
			c.skip(chunk_size);
		}
	}

	// ---- key / values ----
	if ((file_flags & LSSF_KeyValues) && c.ok) {
		const uint32_t n = c.get<uint32_t>();
		for (uint32_t i = 0; i < n && c.ok; ++i)
			c.str(), c.str();
	}

	MarkBulkEdit();
	if (!c.ok)
		return refuse("truncated file");
	for (const NodeRef &ref : nodes_to_disable)
		DisableNode(ref);

	// "Ready Matrices" (scene_load_binary.cpp:888-896): the loaded transforms get their world matrices at once — here one whole-hierarchy
	// upload and one level-ordered pass on the device
	ReadyWorldMatrices();
	ComputeWorldMatrices();
	return true;
}

bool LoadSceneBinaryFromMemory(const void *data, size_t size, const char *name, Scene &scene, PipelineResources &resources, const PipelineInfo &pipeline,
	LoadSceneContext &ctx, uint32_t flags) {
	return scene.Load_binary(data, size, name, resources, pipeline, ctx, flags);
}

bool LoadSceneBinaryFromFile(const char *path, Scene &scene, PipelineResources &resources, const PipelineInfo &pipeline, LoadSceneContext &ctx, uint32_t flags,
	bool) {
	FILE *f = fopen(path, "rb");
	if (!f) {
		char msg[256];
		snprintf(msg, sizeof(msg), "Cannot load scene '%s', invalid read handle", path);
		warn(msg);
		return false;
	}
	std::vector<uint8_t> buf;
	fseek(f, 0, SEEK_END);
	const long n = ftell(f);
	fseek(f, 0, SEEK_SET);
	buf.resize(n > 0 ? size_t(n) : 0);
	const size_t got = buf.empty() ? 0 : fread(buf.data(), 1, buf.size(), f);
	fclose(f);
	return scene.Load_binary(buf.data(), got, path, resources, pipeline, ctx, flags);
}

} // namespace hg
