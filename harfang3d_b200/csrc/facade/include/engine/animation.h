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

struct Anim {
	std::vector<AnimTrackT<bool>> bool_tracks;           // "Enable" (not on the metric path)
	std::vector<AnimTrackHermiteT<float>> float_tracks;   // "Camera.Fov": one value per camera, evaluated on the host (EvaluateHermite, animation.h:155-169)
	std::vector<AnimTrackHermiteT<Vec3>> vec3_tracks;     // "Position", "Rotation" (Euler), "Scale"
	std::vector<AnimTrackT<Quaternion>> quat_tracks;     // "Rotation" when AF_UseQuaternionForRotation
	time_ns t_start{0}, t_end{0};
	uint8_t flags{0};
};

template <typename Track> void SortAnimTrackKeys(Track &track);

} // namespace hg
