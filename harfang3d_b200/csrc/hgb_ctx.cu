// Host side of the C ABI (include/hgb200.h): owns the device-resident, depth-sorted SoA mirror of one hg::Scene (or one batch of
// scene instances), rebuilds it when the hierarchy is edited, and enqueues the per-frame kernels. No CPU fallback exists:
// every entry point that computes needs the CUDA device the context was created on.
#include "../../include/hgb200.h"
#include "hgb_kernels2.cuh"
#include "hgb_kernels_tma.cuh"
#include "hgb_kernels_peer.cuh"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <unordered_map>
#include <array>

using namespace hgb;

static thread_local std::string g_err;

static int fail(const char *fmt, ...) {
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	g_err = buf;
	return 1;
}

// error text of entry points that live in other translation units of the library (hgb_noalloc.cu); not exported
extern "C" __attribute__((visibility("hidden"))) void hgbi_set_error(const char *msg) { g_err = msg ? msg : ""; }

#define CK(call)                                                                                                                                               \
	do {                                                                                                                                                       \
		cudaError_t e_ = (call);                                                                                                                               \
		if (e_ != cudaSuccess)                                                                                                                                 \
			return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);                                                           \
	} while (0)

namespace {

constexpr uint32_t kClockRing = 8;

// device buffer that only grows; contents are NOT preserved across a grow
struct DBuf {
	void *p = nullptr;
	size_t cap = 0;
	cudaError_t reserve(size_t bytes) {
		if (bytes <= cap)
			return cudaSuccess;
		if (p)
			cudaFree(p);
		p = nullptr;
		cap = 0;
		const size_t want = (bytes + 255) & ~size_t(255);
		cudaError_t e = cudaMalloc(&p, want);
		if (e == cudaSuccess)
			cap = want;
		return e;
	}
	void release() {
		if (p)
			cudaFree(p);
		p = nullptr;
		cap = 0;
	}
	template <typename T> T *as() const { return static_cast<T *>(p); }
};

struct HBuf { // pinned host staging, grows
	void *p = nullptr;
	size_t cap = 0;
	cudaError_t reserve(size_t bytes) {
		if (bytes <= cap)
			return cudaSuccess;
		if (p)
			cudaFreeHost(p);
		p = nullptr;
		cap = 0;
		cudaError_t e = cudaMallocHost(&p, bytes);
		if (e == cudaSuccess)
			cap = bytes;
		return e;
	}
	void release() {
		if (p)
			cudaFreeHost(p);
		p = nullptr;
		cap = 0;
	}
	template <typename T> T *as() const { return static_cast<T *>(p); }
};

// per-kernel CUDA-event timing (hgb_set_profiling): every launch of a profiled frame is bracketed by an event pair on the
// context's stream; classes follow HGB_K_* of the header
struct Prof {
	bool on = false;
	std::vector<cudaEvent_t> ev; // pairs
	std::vector<int> cls;        // class of pair i
	size_t used = 0;             // pairs in flight
	double ms[HGB_K_COUNT] = {};
	uint64_t n[HGB_K_COUNT] = {};
};

struct EntryTable {
	uint32_t n = 0;
	std::vector<uint32_t> mtx_slot, bounds_idx;
	std::vector<uint8_t> flags;
	std::vector<uint16_t> view;
	DBuf mtx_pos, meta, view_d;      // per entry (entry order)
	DBuf vis[2];                     // per entry visibility byte, one buffer per world-buffer parity: the tail of frame N (compaction,
	                                 // gather) may still read its bytes while the level kernels of frame N + 1 write theirs
	DBuf out[kMaxViews];             // surviving indices per view
	DBuf keys[kMaxViews];            // sort keys of the survivors (hgb_sort_keys)
	DBuf counts;                     // kMaxViews uint32
	DBuf records, records_out;       // optional: one opaque 16-byte record per entry (hgb_set_entry_records) and the gathered survivors
	uint32_t n_records = 0;
	HBuf records_h;                  // pinned landing zone of hgb_get_visible_records_ptr
	cudaEvent_t rec_count_ev = nullptr; // pipelined fetch: count known / chunk i landed
	std::vector<cudaEvent_t> rec_ev;
	uint32_t rec_chunks = 0;
	uint32_t last_count[kMaxViews] = {}; // what the host last read: sizes the next one-round-trip fetch (fetch_visible)
	DBuf ticket, status;             // compaction scratch (per view slices)
	uint32_t n_tiles = 0;
	uint32_t n_dev = 0;              // entries the device tables and the compaction cover: n at the last table build + the append slack
	                                 // (entries [n, n_dev) are disabled and their visibility bytes stay zero)
};

// peer-memory result gather (hgb_peer_init / hgb_peer_connect, hgb_kernels_peer.cuh)
struct Peer {
	bool on = false;
	uint32_t rank = 0, world = 0, dst = 0;
	uint32_t cap[kMaxPeers] = {}, aux_n[kMaxPeers] = {}, entry_base[kMaxPeers] = {}, list_off[kMaxPeers] = {}, aux_off[kMaxPeers] = {};
	uint32_t aux_base[kMaxPeers + 1] = {};
	size_t half_words = 0;
	void *mailbox = nullptr;      // own: ack, step, error
	void *inbox_local = nullptr;  // destination only
	uint32_t *inbox = nullptr;    // the destination's inbox as seen from this rank (peer mapping or local)
	uint32_t *ack_ptr[kMaxPeers] = {}; // destination only: every rank's ack word
	std::vector<void *> opened;
};

} // namespace

struct hgb_ctx {
	int device = 0;
	bool use_graph = true;
	cudaStream_t own_stream = nullptr, stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	uint64_t launches = 0;
	bool in_post_phase = false;   // enqueue_frame_kernels is recording the frame's asynchronous tail
	uint32_t compact_ctas = 148;  // persistent CTAs of the tail's compaction (HGB_COMPACT_CTAS; 0 = one CTA per tile)
	uint64_t rebuilds = 0; // depth sorts + table rebuilds + graph captures so far (hgb_rebuild_count)
	Prof prof;

	// topology (host)
	uint32_t n_slots = 0, n_pos = 0; // n_pos = used slots (positions [0,n_pos) are live; holes get positions >= n_pos)
	std::vector<uint32_t> pos_of_slot, slot_of_pos, parent_pos_h;
	std::vector<uint32_t> stamp_w, stamp_r; // hgb_set_matrices: run stamps per position
	std::vector<uint32_t> level_off; // level d = positions [level_off[d], level_off[d+1])
	bool identity = true;
	bool entries_dirty = true; // CSR by position must be rebuilt

	// device scene
	DBuf tiles, world[2], preset, pos_of_slot_d; // tiles: node tiles (TRS rows, parent, first entry), see hgb_kernels.cuh
	uint32_t n_tiles = 0;
	bool transpose_store = true;
	bool tma_level = false;   // A/B switch: persistent TMA-staged level kernel instead of one thread per node with LDG rows (measured slower: r01c)
	uint64_t top_budget = 500000;  // k_update_top covers the leading levels whose recomputed work (sum of nodes x depth) stays below this; 0 = one launch per level
	uint32_t level_block = 128;    // threads per CTA of the level kernel: small CTAs recycle their slots sooner (HGB_LEVEL_BLOCK=256: A/B, +2.5 % frame time)
	bool multi_compact = true;     // all views of an entry table compacted in one pass (HGB_MULTI_COMPACT=0: one k_compact launch per view; A/B)
	int level_minb = 6;            // resident CTAs per SM the level kernel's registers are held to (HGB_LEVEL_MINB=5: 48 registers; A/B)
	bool overlap_sampler = false; // opt-in: sample tracks by target level on a second stream, overlapped with the level kernels (measured slower: r01c)
	cudaStream_t stream2 = nullptr;
	std::vector<cudaEvent_t> ev_chunk; // [0] fork, [1] join, [2+g] chunk of level group g
	std::vector<uint32_t> trk_pos_h;   // target position of every round-0 track, ascending (chunk boundaries of the overlapped sampler)
	// in-place edits (hgb_options.append_slack > 0): free positions behind the last level, free entries behind each table
	uint32_t append_slack = 0;
	uint32_t cap_slots = 0;            // device arrays hold this many slots / positions (n_slots at the last depth sort + append_slack)
	std::vector<uint8_t> live_h;       // per position: computed node (sorted region or append region)
	std::vector<uint8_t> late_depth_h; // per position of the append region: ancestors inside the region (chain length of k_update_top)
	std::vector<uint32_t> child_cnt_h; // per position: live children
	std::vector<uint8_t> has_entry_h;  // per position: rows 10 / 11 hold an entry
	DBuf edit_d;                       // scratch of the in-place edit kernels
	// removed transforms: entries [0, tomb[slot][kind]) of that slot's display entries belonged to the node that was removed and stay
	// disabled for good, also when the slot is handed out again and the tables are rebuilt from the host arrays
	std::unordered_map<uint32_t, std::array<uint32_t, 2>> tomb;
	uint32_t slice_nodes = 0;          // >0: with the overlapped sampler, big levels are cut into slices of this many positions (multiple of 256)
	uint32_t persistent_ctas = 148 * 6; // resident CTAs of the device for the persistent level kernel
	uint32_t lookahead = 0; // positions (see SceneDev::lookahead)
	int cur = 0; // world[cur] = current frame
	bool preset_pending = false, preset_dirty = false;
	DBuf preset_pos, preset_mat;
	uint32_t n_preset = 0;

	DBuf bounds;
	std::vector<float> bounds_stage; // interleaved host copy during an upload
	uint32_t n_bounds = 0;
	EntryTable ent[2];
	DBuf ent_off, pent, pos_view; // overflow CSR of nodes with several entries; per-position instance id

	Views views{};          // planes -> view_planes (device)
	DBuf view_planes, view_planes_x; // plain {x,y,z,-2w} and expanded {x,|x|,y,|y|} {z,|z|,-2w,0} forms
	bool instance_mode = false;
	uint32_t n_inst = 0;
	DBuf inst_planes, inst_planes_x, inst_start, inst_rank;

	// skins
	uint32_t n_bones = 0;
	DBuf bone_mtx_pos, bone_bind_idx, bind_pose, bone_out[2];
	std::vector<uint32_t> bone_mtx_slot;

	// scene / instance bounding boxes (hgb_set_objects, hgb_minmax)
	std::vector<uint32_t> obj_slot_h, obj_ob_h, obj_seg_h;
	uint32_t n_obj = 0, n_obounds = 0, n_seg = 0;
	bool obj_dirty = true;
	DBuf obj_pos, obj_meta, obj_seg, obounds, mm_acc;

	// "Enable" bool tracks (hgb_set_enable_tracks): evaluated on the device, state folded back into the host tables before any rebuild
	std::vector<hgb_track> en_tracks_h;      // evaluation order
	std::vector<uint32_t> en_target_slot_h;  // unique target slots, ascending
	std::vector<uint8_t> en_disabled_h;      // last known state per target (valid after a fold)
	std::vector<uint8_t> en_known_h;         // 1 once a fold has seen the target
	uint32_t n_en_keys = 0;
	bool en_state_dirty = false;             // the device sampled enable tracks since the last fold
	bool en_resolved = false;                // device tables below match the current layout
	DBuf en_target_pos, en_off, en_tracks_d, en_keys_d, en_scratch;

	Peer peer;
	DBuf peer_out_list, peer_rank_off, peer_aux_out;

	// animation
	DBuf keys, cursor, tstate, tracks_d, play_t_d; // keys: 32-byte records, vec3 keys then quaternion keys
	DBuf key_rec;                          // 128-byte interval records, one per key (k_build_key_records)
	bool records_dirty = true;
	uint32_t n_v3_keys = 0, n_q_keys = 0;
	DBuf anim_val, bind_pos_d, bind_group_d; // track-driven (node, pos|rot|scale) triples: values + where they are bound
	uint32_t n_bound = 0;                    // triples currently bound on the device (tiles hold slots, not values)
	uint32_t n_tracks = 0, n_play = 0;
	std::vector<hgb_track> tracks_h;
	std::vector<uint32_t> round_off; // tracks sorted so that tracks sharing a target lane fall in successive rounds
	HBuf play_t_h; // kClockRing slots of n_play clocks
	cudaEvent_t clock_ev[8] = {};
	bool clock_used[8] = {};
	uint32_t clock_slot = 0;
	HBuf count_h, visible_h[2]; // pinned landing zones of hgb_get_visible / hgb_get_visible_ptr

	HBuf staging;  // pinned host staging for uploads/downloads
	DBuf staging_d; // device staging (slot-order data before scatter / after gather)
	DBuf flush;
	DBuf adhoc_pos, adhoc_meta, adhoc_bounds, adhoc_planes, adhoc_vis, adhoc_out, adhoc_scratch; // hgb_cull_list

	// CUDA graph of one frame
	cudaGraphExec_t graph_exec[2][2] = {}; // [fused][world-buffer parity]
	bool graph_valid[2][2] = {};
	uint64_t graph_kernels[2][2] = {};
	// Asynchronous tail (hgb_options.async_post): the frame is two graphs — main (track sampling + levels) on the context's stream and
	// post (compaction + the peer gather) on stream2 — and the post graph of frame N overlaps the main graph of frame N + 1: the two
	// touch different visibility buffers (EntryTable::vis[parity]). Frame N + 2 reuses the buffers of frame N and waits for its tail;
	// every other entry point joins the pending tails first (join_post).
	bool pdl_levels = true; // level d+1 launched with a programmatic dependency on level d (HGB_PDL=0 in the environment: plain, A/B)
	bool pdl_first = true; // the first level is chained to the track sampler (HGB_PDL_FIRST=0: plain launch; A/B)
	bool pdl_late_wait = true; // see SceneDev::late_wait (HGB_PDL_LATE=0: wait ahead of the sampled values, sampler -> first level chained instead; A/B)
	bool async_post = true;
	// ... chosen per frame when nobody forced it: a caller that reads the frame's result on the host after EVERY update (a synchronising
	// getter between two updates) leaves the tail nothing to overlap with, and the one-graph frame — tail chained to the last level by a
	// programmatic dependency, no cross-stream event — ends sooner. Two such frames in a row switch to it, the first update that finds the
	// previous result unread switches back.
	bool async_auto = true;
	bool result_read = false;     // a result getter synchronised since the last hgb_update
	uint32_t sync_streak = 0;
	bool last_frame_async = false;
	cudaGraphExec_t phase_exec[3][2] = {}; // [phase][world-buffer parity]
	bool phase_valid[3][2] = {};
	uint64_t phase_kernels[3][2] = {};
	cudaEvent_t ev_lvl_done = nullptr, ev_post_done[2] = {};
	bool post_pending[2] = {};
};

// the asynchronous tail of the last frame must be over before anything else touches the context's stream or tables
static void join_post(hgb_ctx *c) {
	if (!c)
		return;
	for (int par = 0; par < 2; ++par)
		if (c->post_pending[par]) {
			cudaSetDevice(c->device);
			cudaStreamWaitEvent(c->stream, c->ev_post_done[par], 0);
			c->post_pending[par] = false;
		}
}

// bracket one kernel launch with events when profiling (never during graph capture: profiled frames use plain launches)
static void prof_begin(hgb_ctx *c, int cls) {
	Prof &p = c->prof;
	if (!p.on)
		return;
	if (p.used * 2 == p.ev.size()) {
		cudaEvent_t a = nullptr, b = nullptr;
		cudaEventCreate(&a);
		cudaEventCreate(&b);
		p.ev.push_back(a);
		p.ev.push_back(b);
		p.cls.push_back(cls);
	}
	p.cls[p.used] = cls;
	cudaEventRecord(p.ev[p.used * 2], c->stream);
}
static void prof_end(hgb_ctx *c) {
	Prof &p = c->prof;
	if (!p.on)
		return;
	cudaEventRecord(p.ev[p.used * 2 + 1], c->stream);
	++p.used;
}
static int prof_collect(hgb_ctx *c) {
	Prof &p = c->prof;
	if (!p.used)
		return 0;
	CK(cudaStreamSynchronize(c->stream));
	for (size_t i = 0; i < p.used; ++i) {
		float ms = 0.f;
		CK(cudaEventElapsedTime(&ms, p.ev[i * 2], p.ev[i * 2 + 1]));
		p.ms[p.cls[i]] += ms;
		++p.n[p.cls[i]];
	}
	p.used = 0;
	return 0;
}

static void invalidate_graph(hgb_ctx *c) {
	for (auto &a : c->graph_valid)
		for (bool &v : a)
			v = false;
	for (auto &a : c->phase_valid)
		for (bool &v : a)
			v = false;
}

extern "C" {

const char *hgb_last_error(void) { return g_err.c_str(); }

int hgb_device_count(void) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess)
		return 0;
	return n;
}

hgb_ctx *hgb_create(int device, const hgb_options *opt) {
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) {
		fail("hgb_create: no CUDA device (%s). This library has no CPU path.", e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
		return nullptr;
	}
	if (device < 0 || device >= n) {
		fail("hgb_create: device %d out of range (%d devices)", device, n);
		return nullptr;
	}
	if ((e = cudaSetDevice(device)) != cudaSuccess) {
		fail("cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
		return nullptr;
	}
	if (const char *e_l2 = getenv("HGB_L2_FETCH")) { // A/B: DRAM -> L2 fetch granularity in bytes (32 / 64 / 128); device-wide limit
		size_t before = 0;
		cudaDeviceGetLimit(&before, cudaLimitMaxL2FetchGranularity);
		const cudaError_t el = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, size_t(atoi(e_l2)));
		size_t after = 0;
		cudaDeviceGetLimit(&after, cudaLimitMaxL2FetchGranularity);
		fprintf(stderr, "hgb: L2 fetch granularity %zu -> %zu (%s)\n", before, after, cudaGetErrorString(el));
	}
	hgb_ctx *c = new hgb_ctx;
	c->device = device;
	c->use_graph = opt ? opt->use_graph != 0 : true;
	{
		cudaDeviceProp prop;
		if (cudaGetDeviceProperties(&prop, device) == cudaSuccess)
		{
			c->lookahead = uint32_t(prop.multiProcessorCount) * 6u * 256u; // one wave of resident CTAs (6 x 256 threads per SM)
			c->persistent_ctas = uint32_t(prop.multiProcessorCount) * 6u;
		}
		if (opt && opt->lookahead != 0)
			c->lookahead = opt->lookahead < 0 ? 0u : uint32_t(opt->lookahead); // A/B switch: -1 = off, n = positions
	}
	c->tma_level = opt ? opt->tma_level != 0 : false;
	if (opt && opt->top_walk != 0)
		c->top_budget = opt->top_walk < 0 ? 0u : uint64_t(opt->top_walk); // A/B switch: -1 = one launch per level, n = budget
	if (const char *e_top = getenv("HGB_TOP_BUDGET"))
		c->top_budget = strtoull(e_top, nullptr, 10);
	if (const char *e_mc = getenv("HGB_MULTI_COMPACT"))
		c->multi_compact = e_mc[0] != '0';
	if (const char *e_blk = getenv("HGB_LEVEL_BLOCK"))
		c->level_block = atoi(e_blk) == 256 ? 256u : (atoi(e_blk) == 64 ? 64u : 128u);
	if (const char *e_minb = getenv("HGB_LEVEL_MINB"))
		c->level_minb = atoi(e_minb) == 5 ? 5 : 6;
	c->overlap_sampler = opt ? opt->overlap_sampler != 0 : false;
	if (opt && opt->slice_nodes > 0)
		c->slice_nodes = (uint32_t(opt->slice_nodes) + 255u) & ~255u;
	if (opt && opt->append_slack > 0)
		c->append_slack = uint32_t(opt->append_slack);
	if (const char *e = getenv("HGB_COMPACT_CTAS"))
		c->compact_ctas = uint32_t(atoi(e));
	else {
		cudaDeviceProp prop{};
		if (cudaGetDeviceProperties(&prop, device) == cudaSuccess)
			c->compact_ctas = uint32_t(prop.multiProcessorCount);
	}
	if (opt && opt->async_post != 0)
		c->async_post = opt->async_post > 0, c->async_auto = false; // 0 = default (chosen per frame), 1 = always, -1 = never
	if (const char *e_ap = getenv("HGB_ASYNC_POST")) // A/B: 0 never, 1 always, 2 chosen per frame
		c->async_post = e_ap[0] != '0', c->async_auto = e_ap[0] == '2';
	if (const char *e_pdl = getenv("HGB_PDL"))
		c->pdl_levels = e_pdl[0] != '0';
	if (const char *e_first = getenv("HGB_PDL_FIRST"))
		c->pdl_first = e_first[0] != '0';
	if (const char *e_late = getenv("HGB_PDL_LATE"))
		c->pdl_late_wait = e_late[0] != '0';
	c->transpose_store = opt ? opt->row_store == 0 : true; // reserved[0] = 1: per-thread 48-byte row stores (A/B switch for profiling)
	if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&c->ev0) != cudaSuccess ||
		cudaEventCreate(&c->ev1) != cudaSuccess) {
		fail("hgb_create: stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
		delete c;
		return nullptr;
	}
	c->stream = c->own_stream;
	int prio_lo = 0, prio_hi = 0;
	cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
	if (cudaEventCreateWithFlags(&c->ev_lvl_done, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&c->ev_post_done[0], cudaEventDisableTiming) != cudaSuccess ||
		cudaEventCreateWithFlags(&c->ev_post_done[1], cudaEventDisableTiming) != cudaSuccess ||
		cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio_hi) != cudaSuccess) { // the tail is short and somebody waits for it: ahead of the next frame's CTAs
		fail("hgb_create: stream creation failed: %s", cudaGetErrorString(cudaGetLastError()));
		delete c;
		return nullptr;
	}
	for (cudaEvent_t &e : c->clock_ev)
		cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
	for (int i = 0; i < 96; ++i) { // fork/join/chunk events of the overlapped sampler, created up front (never during graph capture)
		cudaEvent_t e = nullptr;
		if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess)
			c->ev_chunk.push_back(e);
	}
	c->views.n_views = 0;
	if (c->view_planes.reserve(kMaxViews * 96) != cudaSuccess || c->view_planes_x.reserve(kMaxViews * 192) != cudaSuccess) {
		fail("hgb_create: out of device memory");
		delete c;
		return nullptr;
	}
	c->views.planes = c->view_planes.as<float4>();
	c->views.planes_x = c->view_planes_x.as<float4>();
	return c;
}

static void peer_release(hgb_ctx *c);

void hgb_destroy(hgb_ctx *c) {
	if (!c)
		return;
	cudaSetDevice(c->device);
	join_post(c);
	cudaStreamSynchronize(c->stream);
	if (c->stream2)
		cudaStreamSynchronize(c->stream2);
	peer_release(c);
	for (auto &a : c->phase_exec)
		for (cudaGraphExec_t g : a)
			if (g)
				cudaGraphExecDestroy(g);
	for (auto &a : c->graph_exec)
		for (cudaGraphExec_t g : a)
			if (g)
				cudaGraphExecDestroy(g);
	DBuf *bufs[] = {&c->tiles, &c->world[0], &c->world[1], &c->preset, &c->pos_of_slot_d, &c->preset_pos, &c->preset_mat, &c->bounds, &c->ent_off,
		&c->pent, &c->pos_view, &c->inst_planes, &c->inst_planes_x, &c->view_planes_x, &c->inst_start, &c->inst_rank, &c->bone_mtx_pos, &c->bone_bind_idx, &c->bind_pose, &c->bone_out[0],
		&c->bone_out[1], &c->keys, &c->cursor, &c->tstate, &c->anim_val, &c->bind_pos_d, &c->bind_group_d, &c->tracks_d, &c->play_t_d, &c->staging_d, &c->flush, &c->adhoc_pos, &c->adhoc_meta, &c->adhoc_bounds, &c->adhoc_planes, &c->adhoc_vis, &c->adhoc_out, &c->adhoc_scratch, &c->view_planes, &c->key_rec,
		&c->en_target_pos, &c->en_off, &c->en_tracks_d, &c->en_keys_d, &c->en_scratch, &c->obj_pos, &c->obj_meta, &c->obj_seg, &c->obounds, &c->mm_acc, &c->peer_out_list, &c->peer_rank_off, &c->peer_aux_out};
	for (DBuf *b : bufs)
		b->release();
	for (auto &t : c->ent) {
		DBuf *eb[] = {&t.mtx_pos, &t.meta, &t.view_d, &t.vis[0], &t.vis[1], &t.counts, &t.ticket, &t.status, &t.records, &t.records_out};
		t.records_h.release();
		if (t.rec_count_ev)
			cudaEventDestroy(t.rec_count_ev);
		for (cudaEvent_t e : t.rec_ev)
			cudaEventDestroy(e);
		for (DBuf *b : eb)
			b->release();
		for (auto &o : t.out)
			o.release();
		for (auto &o : t.keys)
			o.release();
	}
	for (cudaEvent_t e : c->prof.ev)
		cudaEventDestroy(e);
	c->staging.release();
	c->play_t_h.release();
	c->count_h.release();
	c->visible_h[0].release();
	c->visible_h[1].release();
	for (cudaEvent_t e : c->clock_ev)
		if (e)
			cudaEventDestroy(e);
	cudaEventDestroy(c->ev0);
	cudaEventDestroy(c->ev1);
	for (cudaEvent_t e : c->ev_chunk)
		cudaEventDestroy(e);
	if (c->stream2)
		cudaStreamDestroy(c->stream2);
	cudaStreamDestroy(c->own_stream);
	delete c;
}

int hgb_set_stream(hgb_ctx *c, void *s) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaStreamSynchronize(c->stream));
	c->stream = s ? static_cast<cudaStream_t>(s) : c->own_stream;
	invalidate_graph(c);
	return 0;
}

int hgb_sync(hgb_ctx *c) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

uint64_t hgb_launch_count(hgb_ctx *c) { return c ? c->launches : 0; }
uint64_t hgb_rebuild_count(hgb_ctx *c) { return c ? c->rebuilds : 0; }
int hgb_frame_was_async(hgb_ctx *c) { return c && c->last_frame_async ? 1 : 0; }
uint32_t hgb_level_count(hgb_ctx *c) { return c && !c->level_off.empty() ? uint32_t(c->level_off.size() - 1) : 0; }
int hgb_is_identity_order(hgb_ctx *c) { return c && c->identity ? 1 : 0; }

int hgb_time_begin(hgb_ctx *c) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaEventRecord(c->ev0, c->stream));
	return 0;
}
int hgb_time_end(hgb_ctx *c, float *ms) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaEventRecord(c->ev1, c->stream));
	CK(cudaEventSynchronize(c->ev1));
	CK(cudaEventElapsedTime(ms, c->ev0, c->ev1));
	return 0;
}
int hgb_set_profiling(hgb_ctx *c, int on) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (prof_collect(c))
		return 1;
	c->prof.on = on != 0;
	return 0;
}
int hgb_get_profile(hgb_ctx *c, double *ms_by_class, uint64_t *launches_by_class, int reset) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (prof_collect(c))
		return 1;
	for (int k = 0; k < HGB_K_COUNT; ++k) {
		if (ms_by_class)
			ms_by_class[k] = c->prof.ms[k];
		if (launches_by_class)
			launches_by_class[k] = c->prof.n[k];
		if (reset) {
			c->prof.ms[k] = 0;
			c->prof.n[k] = 0;
		}
	}
	return 0;
}
int hgb_flush_l2(hgb_ctx *c, size_t bytes) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(c->flush.reserve(bytes));
	CK(cudaMemsetAsync(c->flush.p, 0x5a, bytes, c->stream));
	return 0;
}

} // extern "C"

#include "hgb_ctx_scene.inl"
#include "hgb_ctx_peer.inl"
#include "hgb_ctx_frame.inl"
