// hg::Scene facade — keyframe containers of the scene-update path (harfang/engine/animation.h:26-52,199-218). Same member names so
// that code filling tracks compiles unchanged; key storage is a std::vector (the reference uses std::deque) because the keys are
// flattened into one device pool at bind time anyway. Evaluation happens on the GPU (k_sample_tracks), not here.
#pragma once
#include "foundation/hg_math_types.h"
#include <string>
#include <vector>

namespace hg {

enum AnimLoopMode : uint8_t { ALM_Once, ALM_Infinite, ALM_Loop };
enum Easing : uint8_t { E_Linear = 0 }; // only linear playback reaches the evaluation: the reference's easing branch is dead code (scene.cpp:2418-2419)

template <typename T> struct AnimKeyT {
	time_ns t;
	T v;
};
template <typename T> struct AnimKeyHermiteT {
	time_ns t;
	T v;
	float tension{0.f}, bias{0.f};
};
template <typename T> struct AnimTrackT {
	std::string target;
	std::vector<AnimKeyT<T>> keys;
};
template <typename T> struct AnimTrackHermiteT {
	std::string target;
	std::vector<AnimKeyHermiteT<T>> keys;
};

static const uint8_t AF_UseQuaternionForRotation = 0x01;
static const uint8_t AF_Instantiated = 0x10; // non-serialized: the anim came with an instantiated scene (animation.h:202)

struct InstanceAnimKey { // animation.h:171-175: which anim of an instantiated scene plays from this key on
	std::string anim_name;
	AnimLoopMode loop_mode{ALM_Once};
	float t_scale{1.f};
};

struct Anim {
	std::vector<AnimTrackT<bool>> bool_tracks;           // "Enable": evaluated on the device (k_sample_enable)
	std::vector<AnimTrackHermiteT<float>> float_tracks;   // "Camera.Fov", "Light.DiffuseIntensity", "Light.SpecularIntensity": single host-side values
	std::vector<AnimTrackHermiteT<Vec3>> vec3_tracks;     // "Position", "Rotation" (Euler), "Scale": evaluated on the device (k_sample_tracks)
	std::vector<AnimTrackHermiteT<Vec4>> vec4_tracks;     // "Material.<slot>.<value>": single host-side values
	std::vector<AnimTrackT<Quaternion>> quat_tracks;     // "Rotation" when AF_UseQuaternionForRotation (device)
	std::vector<AnimTrackHermiteT<Color>> color_tracks;   // "Light.Diffuse", "Light.Specular": single host-side values
	AnimTrackT<InstanceAnimKey> instance_anim_track;      // step track driving the anims of the node's instantiated scene (scene.cpp:2217-2268)
	time_ns t_start{0}, t_end{0};
	uint8_t flags{0};
};

template <typename Track> void SortAnimTrackKeys(Track &track);

} // namespace hg
