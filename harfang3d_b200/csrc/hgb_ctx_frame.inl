// Per-frame half of the C ABI: enqueue (or replay as a CUDA graph) the kernels of one Scene::Update (+ fused cull), stand-alone
// cull, bone matrices, and the device->host getters. Included by hgb_ctx.cu.

static SceneDev scene_dev(hgb_ctx *c, bool with_preset) {
	SceneDev s{};
	s.tiles = c->tiles.as<uint32_t>();
	s.anim_val = c->anim_val.as<float4>();
	s.transpose_store = c->transpose_store ? 1 : 0;
	s.lookahead = c->lookahead;
	s.late_wait = c->pdl_late_wait ? 1 : 0;
	s.world = c->world[c->cur].as<float>();
	s.preset = with_preset ? c->preset.as<uint8_t>() : nullptr;
	s.ent_off = c->ent_off.as<uint32_t>();
	s.pent = c->pent.as<uint2>();
	s.bounds6 = c->bounds.as<float>();
	s.vis0 = c->ent[0].vis[c->cur].as<uint8_t>();
	s.vis1 = c->ent[1].vis[c->cur].as<uint8_t>();
	s.inst_planes = c->inst_planes.as<float4>();
	s.inst_planes_x = c->inst_planes_x.as<float4>();
	s.pos_view = c->pos_view.as<uint16_t>();
	return s;
}

// Launch on the context's stream; with `pdl` as a programmatic dependent of the kernel enqueued just before it: the grid may be
// scheduled while that kernel drains, and its griddepcontrol.wait holds it until the predecessor's writes are visible. Only for
// kernels that execute griddepcontrol.wait before they touch anything the predecessor (or, transitively, its predecessors) wrote.
template <typename... P, typename... A> static void launch_after(hgb_ctx *c, bool pdl, void (*kernel)(P...), dim3 grid, dim3 block, A... args) {
	cudaLaunchConfig_t cfg{};
	cfg.gridDim = grid;
	cfg.blockDim = block;
	cfg.stream = c->stream;
	cudaLaunchAttribute at[1];
	at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	at[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = at;
	cfg.numAttrs = pdl ? 1 : 0;
	cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}

// ticket and look-back state of every compaction of the frame (enqueued ahead of the level kernels when the compaction is chained to
// the last level by a programmatic dependency: a memset node between the two kernels would turn that edge into a full one)
static int enqueue_compaction_reset(hgb_ctx *c) {
	const int n_views = c->instance_mode ? 1 : c->views.n_views;
	for (int k = 0; k < 2; ++k) {
		EntryTable &t = c->ent[k];
		if (!t.n_dev)
			continue;
		CK(cudaMemsetAsync(t.ticket.p, 0, kMaxViews * 4, c->stream));
		CK(cudaMemsetAsync(t.status.p, 0, size_t(t.n_tiles) * (n_views > 1 ? kMaxViews : 1) * 4, c->stream)); // the one-pass kernel walks 2 / 4 / 8 chains
	}
	return 0;
}

// compaction of every (kind, view) list from the vis bytes; returns the number of kernels enqueued through *n_k. `chained`: the reset
// went out earlier and the kernel enqueued just before this call signals its dependents (a level kernel)
static int enqueue_compaction(hgb_ctx *c, uint64_t *n_k, bool with_peer = false, bool chained = false) {
	const int n_views = c->instance_mode ? 1 : c->views.n_views;
	with_peer = with_peer && c->peer.on;
	if (!chained && enqueue_compaction_reset(c))
		return 1;
	for (int k = 0; k < 2; ++k) {
		EntryTable &t = c->ent[k];
		if (!t.n_dev)
			continue;
		const bool peer_here = with_peer && k == 0;
		if (n_views > 1 && !peer_here && c->multi_compact) { // every view of the table in one pass over the visibility bytes
			CompactMultiDev d{};
			d.vis = t.vis[c->cur].as<uint8_t>();
			d.n = t.n_dev;
			for (int v = 0; v < kMaxViews; ++v)
				d.out[v] = v < n_views ? t.out[v].as<uint32_t>() : nullptr;
			d.count = t.counts.as<uint32_t>();
			d.ticket = t.ticket.as<uint32_t>();
			d.status = t.status.as<uint32_t>();
			d.n_tiles = t.n_tiles;
			prof_begin(c, HGB_K_COMPACT);
			switch (n_views) { // exactly the views there are: an empty view still costs its scans and barriers
			case 2: launch_after(c, chained, k_compact_multi<2>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			case 3: launch_after(c, chained, k_compact_multi<3>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			case 4: launch_after(c, chained, k_compact_multi<4>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			case 5: launch_after(c, chained, k_compact_multi<5>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			case 6: launch_after(c, chained, k_compact_multi<6>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			case 7: launch_after(c, chained, k_compact_multi<7>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			default: launch_after(c, chained, k_compact_multi<8>, dim3(t.n_tiles), dim3(kCompactThreads), d); break;
			}
			prof_end(c);
			++*n_k;
			continue;
		}
		for (int v = 0; v < n_views; ++v) {
			CompactDev d{};
			d.vis = t.vis[c->cur].as<uint8_t>();
			d.n = t.n_dev;
			d.bit = uint32_t(v);
			d.out = t.out[v].as<uint32_t>();
			d.count = t.counts.as<uint32_t>() + v;
			d.ticket = t.ticket.as<uint32_t>() + v;
			d.status = t.status.as<uint32_t>() + size_t(v) * t.n_tiles;
			if (with_peer && k == 0 && v == 0) { // the visible list leaves for the gather destination from the kernel that produces it
				if (t.n_dev > c->peer.cap[c->peer.rank])
					return fail("peer gather: %u display entries but a region of %u was declared", t.n_dev, c->peer.cap[c->peer.rank]);
				d.out2 = c->peer.inbox + c->peer.list_off[c->peer.rank];
				d.add2 = c->peer.entry_base[c->peer.rank];
				d.out2_half = uint32_t(c->peer.half_words);
				d.out2_step = static_cast<const uint32_t *>(c->peer.mailbox) + 1;
			}
			prof_begin(c, HGB_K_COMPACT);
			// hidden behind the next frame (asynchronous tail): one persistent CTA per SM, so that the level kernels keep the SMs
			// (only with the peer gather, whose stores the fabric can hold up: measured at N = 4 0.488 -> 0.480 ms; alone on a GPU the
			// full grid is faster, 0.461 vs 0.469 ms — profiles/README.md r02)
			const uint32_t grid = (c->in_post_phase && c->peer.on && c->compact_ctas) ? std::min(t.n_tiles, c->compact_ctas) : t.n_tiles;
			launch_after(c, chained, k_compact, dim3(grid), dim3(kCompactThreads), d);
			prof_end(c);
			++*n_k;
		}
		if (c->instance_mode && k == 0 && c->n_inst) {
			prof_begin(c, HGB_K_SEGMENT_RANKS);
			launch_after(c, chained, k_segment_ranks, grid_for(c->n_inst * 32), dim3(256), t.vis[c->cur].as<uint8_t>(), 0u, c->inst_start.as<uint32_t>(), c->n_inst,
				t.status.as<uint32_t>(), c->inst_rank.as<uint32_t>());
			prof_end(c);
			++*n_k;
		}
	}
	CK(cudaGetLastError());
	if (with_peer)
		return peer_enqueue_end(c, n_k);
	return 0;
}

template <int MINB> static void launch_level_v(hgb_ctx *c, const SceneDev &s, bool do_fused, uint32_t b, uint32_t e, bool pdl) {
	const dim3 blk(c->level_block), g = grid_for(e - b, c->level_block);
	if (!do_fused)
		launch_after(c, pdl, k_update_level<false, false, MINB>, g, blk, s, c->views, b, e);
	else if (c->instance_mode)
		launch_after(c, pdl, k_update_level<true, true, MINB>, g, blk, s, c->views, b, e);
	else
		launch_after(c, pdl, k_update_level<true, false, MINB>, g, blk, s, c->views, b, e);
}

static void launch_level(hgb_ctx *c, const SceneDev &s, bool do_fused, uint32_t b, uint32_t e, bool pdl = false) {
	if (c->tma_level) { // persistent warps, TMA-staged tiles (hgb_kernels_tma.cuh)
		const uint32_t n_lt = ((e - 1u) >> 5) - (b >> 5) + 1u;
		const dim3 g(std::min((n_lt + kTmaWarps - 1u) / kTmaWarps, c->persistent_ctas));
		if (!do_fused)
			k_update_level_tma<false, false><<<g, kTmaWarps * 32, 0, c->stream>>>(s, c->views, b, e);
		else if (c->instance_mode)
			k_update_level_tma<true, true><<<g, kTmaWarps * 32, 0, c->stream>>>(s, c->views, b, e);
		else
			k_update_level_tma<true, false><<<g, kTmaWarps * 32, 0, c->stream>>>(s, c->views, b, e);
	} else if (c->level_minb == 5) {
		launch_level_v<5>(c, s, do_fused, b, e, pdl);
	} else {
		launch_level_v<6>(c, s, do_fused, b, e, pdl);
	}
}

// the first `top_levels` levels in one launch (k_update_top): positions [0, e); or the append region [b, e) behind the last level,
// whose chains stop at the first ancestor below b
static void launch_top(hgb_ctx *c, const SceneDev &s, bool do_fused, uint32_t b, uint32_t e, bool pdl) {
	const dim3 g = grid_for(e - b);
	if (!do_fused)
		launch_after(c, pdl, k_update_top<false, false>, g, dim3(256), s, c->views, b, e, b);
	else if (c->instance_mode)
		launch_after(c, pdl, k_update_top<true, true>, g, dim3(256), s, c->views, b, e, b);
	else
		launch_after(c, pdl, k_update_top<true, false>, g, dim3(256), s, c->views, b, e, b);
}

// How many leading levels k_update_top covers: as many as keep the recomputed work (every node of level d evaluates d + 1 local
// matrices) inside the budget, at most kMaxWalk. 1 = the first level alone (no recomputation).
static size_t plan_top_levels(const hgb_ctx *c) {
	const size_t n_levels = c->level_off.size() - 1;
	if (c->top_budget == 0 || c->tma_level)
		return 0;
	uint64_t evals = 0;
	size_t d = 0;
	while (d < n_levels && d < size_t(kMaxWalk)) {
		const uint64_t add = uint64_t(c->level_off[d + 1] - c->level_off[d]) * (d + 1);
		if (d >= 1 && evals + add > c->top_budget)
			break;
		evals += add;
		++d;
	}
	return d;
}

// One frame. Level groups (a level, or a run of small levels) go out in order on the context's stream. The track sampler is HBM-bound
// and issue-light, the level kernels issue-heavy: with `overlap_sampler` the tracks are sampled in chunks by TARGET LEVEL on a second
// stream, each level group waiting only for its own chunk — so the chunk of the (huge) last level streams its keys while the upper
// levels compute. Tracks are sorted by target position (prepare_tables), so a chunk is a contiguous track range.
constexpr int kPhasePre = 1, kPhaseLevels = 2, kPhasePost = 4, kPhaseAll = 7;
static int enqueue_frame_kernels(hgb_ctx *c, bool fused, bool with_preset, uint64_t *n_k, int phases = kPhaseAll) {
	const SceneDev s = scene_dev(c, with_preset);
	const bool do_fused = fused && (c->ent[0].n_dev || c->ent[1].n_dev);
	const size_t n_levels = c->level_off.size() - 1;
	// a rank without display entries still takes part in the gather (it publishes an empty list)
	const bool do_post = do_fused || (fused && c->peer.on);
	// The gather's step is counted by the first kernel of whichever graph holds the frame's tail: all peer kernels of a frame then sit
	// in ONE stream-ordered sequence (the whole frame, or the tail on stream2), and none of them can see another frame's step.
	if (c->peer.on && fused && (phases & kPhasePost))
		peer_enqueue_begin(c, n_k);
	if (phases == kPhasePost) {
		c->in_post_phase = true;
		const int rc = do_post ? enqueue_compaction(c, n_k, true) : 0;
		c->in_post_phase = false;
		return rc;
	}
	const bool has_anim = c->n_tracks && c->n_play;
	AnimDev a{};
	if (has_anim) {
		a.tracks = c->tracks_d.as<Track>();
		a.play_t = c->play_t_d.as<int64_t>();
		a.keys = c->keys.as<uint4>();
		a.rec = c->key_rec.as<uint4>();
		a.quat_base = c->n_v3_keys;
		a.cursor = c->cursor.as<uint32_t>();
		a.tstate = c->tstate.as<uint4>();
		a.anim_val = c->anim_val.as<float4>();
	}
	// level groups: a level, a run of small levels, or — with the overlapped sampler — a slice of a big level, so that the sampler chunk
	// of slice k+1 streams its keys while slice k computes (the level kernel is issue-bound and leaves HBM headroom, the sampler is
	// HBM-bound and leaves issue slots)
	struct Group {
		size_t d0, d1;   // levels [d0, d1): more than one = the top of the hierarchy in one launch (k_update_top)
		uint32_t b, e;   // positions (a slice when d1 == d0 + 1)
		bool top;        // k_update_top: the leading levels, or the append region behind the last level
	};
	const bool want_overlap = has_anim && c->overlap_sampler && !c->prof.on && c->round_off.size() == 2 && !c->trk_pos_h.empty() && phases == kPhaseAll;
	std::vector<Group> groups;
	const size_t n_top = want_overlap ? 0 : plan_top_levels(c);
	if (n_top >= 1 && c->level_off[n_top] > 0)
		groups.push_back({0, n_top, 0u, c->level_off[n_top], true});
	for (size_t d = n_top; d < n_levels; ++d) {
		if (c->level_off[d + 1] == c->level_off[d])
			continue;
		const uint32_t lb = c->level_off[d], le = c->level_off[d + 1];
		const uint32_t slice = want_overlap && c->slice_nodes ? c->slice_nodes : 0xffffff00u;
		for (uint32_t p0 = lb; p0 < le;) {
			const uint32_t p1 = (le - p0 <= slice + slice / 2) ? le : p0 + slice; // slice is a multiple of 256: warps keep their tile alignment
			groups.push_back({d, d + 1, p0, p1, false});
			p0 = p1;
		}
	}
	if (c->append_slack && c->cap_slots > c->n_pos) // nodes appended since the last depth sort (and the free positions among them)
		groups.push_back({n_levels, n_levels + 1, c->n_pos, c->cap_slots, true});

	while (want_overlap && c->ev_chunk.size() < groups.size() + 2) { // 96 exist from hgb_create; more only for very deep or very large scenes
		cudaEvent_t e = nullptr;
		CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
		c->ev_chunk.push_back(e);
	}
	const bool overlap = want_overlap;
	std::vector<char> has_chunk(groups.size(), 0);
	// programmatic dependencies (kernel -> kernel, nothing enqueued in between): sampler -> first level -> ... -> last level -> compactions
	const bool chain = c->pdl_levels && !c->prof.on && !c->tma_level && !overlap;
	const bool chain_post = chain && do_fused && phases == kPhaseAll && !groups.empty();
	if (chain_post && enqueue_compaction_reset(c))
		return 1;
	bool prev_signals = false; // the kernel enqueued last executes griddepcontrol.launch_dependents
	bool prev_is_sampler = false;
	if (has_anim && !overlap && (phases & kPhasePre)) { // sequential: one launch per write round (normally one)
		for (size_t r = 0; r + 1 < c->round_off.size(); ++r) {
			const uint32_t first = c->round_off[r], cnt = c->round_off[r + 1] - first;
			if (!cnt)
				continue;
			prof_begin(c, HGB_K_SAMPLE_TRACKS);
			k_sample_tracks<<<grid_for(cnt, kSampleWarps * 32), kSampleWarps * 32, 0, c->stream>>>(a, first, cnt);
			prof_end(c);
			++*n_k;
			// with the late wait a level reads sampled values ahead of its wait: the level chained to the sampler waits first thing and
			// releases its own dependents only behind that wait (late_wait = 2)
			prev_signals = !c->pdl_late_wait || c->pdl_first;
			prev_is_sampler = true;
		}
	} else if (overlap) {
		CK(cudaEventRecord(c->ev_chunk[0], c->stream)); // fork (also orders the chunks behind this frame's play-clock upload)
		CK(cudaStreamWaitEvent(c->stream2, c->ev_chunk[0], 0));
		for (size_t g = 0; g < groups.size(); ++g) {
			// tracks are sorted by target position: the chunk of positions [b, e) is a contiguous track range
			const uint32_t lo = uint32_t(std::lower_bound(c->trk_pos_h.begin(), c->trk_pos_h.end(), groups[g].b) - c->trk_pos_h.begin());
			const uint32_t hi = uint32_t(std::lower_bound(c->trk_pos_h.begin(), c->trk_pos_h.end(), groups[g].e) - c->trk_pos_h.begin());
			if (hi > lo) {
				k_sample_tracks<<<grid_for(hi - lo, kSampleWarps * 32), kSampleWarps * 32, 0, c->stream2>>>(a, lo, hi - lo);
				++*n_k;
				CK(cudaEventRecord(c->ev_chunk[g + 2], c->stream2));
				has_chunk[g] = 1;
			}
		}
		CK(cudaEventRecord(c->ev_chunk[1], c->stream2)); // join point
	}

	if (!c->en_target_slot_h.empty() && c->n_play && (phases & kPhasePre)) { // "Enable" tracks: node / entry flags of this frame, before the level kernels read them
		EnableDev e{};
		e.target_pos = c->en_target_pos.as<uint32_t>();
		e.off = c->en_off.as<uint32_t>();
		e.tracks = c->en_tracks_d.as<uint4>();
		e.keys = c->en_keys_d.as<uint4>();
		e.play_t = c->play_t_d.as<int64_t>();
		e.tiles = c->tiles.as<uint32_t>();
		e.ent_meta0 = c->ent[0].meta.as<uint32_t>();
		e.ent_meta1 = c->ent[1].meta.as<uint32_t>();
		e.ent_off = c->ent_off.as<uint32_t>();
		e.pent = c->pent.as<uint2>();
		e.n = uint32_t(c->en_target_slot_h.size());
		k_sample_enable<<<grid_for(e.n, 128), 128, 0, c->stream>>>(e);
		++*n_k;
		prev_signals = prev_is_sampler = false; // writes the flag rows the level kernels read before their wait: a full dependency
	}

	if (!(phases & kPhaseLevels)) {
		CK(cudaGetLastError());
		return 0;
	}
	for (size_t g = 0; g < groups.size(); ++g) {
		if (has_chunk[g])
			CK(cudaStreamWaitEvent(c->stream, c->ev_chunk[g + 2], 0));
		prof_begin(c, HGB_K_UPDATE_LEVEL);
		if (groups[g].top) {
			// waits first thing and releases its dependents behind the wait (it reads sampled values throughout)
			launch_top(c, s, do_fused, groups[g].b, groups[g].e, chain && prev_signals);
			prev_signals = true;
			prev_is_sampler = false;
		} else {
			// a level that directly follows the sampler or another level may start while that one drains (k_update_level waits before
			// it reads sampled values or parent rows)
			SceneDev sl = s;
			if (chain && prev_signals && prev_is_sampler && c->pdl_late_wait)
				sl.late_wait = 2;
			launch_level(c, sl, do_fused, groups[g].b, groups[g].e, chain && prev_signals);
			prev_signals = true;
			prev_is_sampler = false;
		}
		prof_end(c);
		++*n_k;
	}
	if (overlap)
		CK(cudaStreamWaitEvent(c->stream, c->ev_chunk[1], 0));
	CK(cudaGetLastError());
	if (do_post && (phases & kPhasePost))
		return enqueue_compaction(c, n_k, true, chain_post && prev_signals);
	return 0;
}

// asynchronous tail: the frame as two graphs; main (pre + levels) on the context's stream, post on stream2 (see hgb_ctx::async_post)
static int launch_phase_graph(hgb_ctx *c, int phase_idx, int phase_mask, cudaStream_t on) {
	const int par = c->cur;
	if (!c->phase_valid[phase_idx][par]) {
		if (c->phase_exec[phase_idx][par]) {
			cudaGraphExecDestroy(c->phase_exec[phase_idx][par]);
			c->phase_exec[phase_idx][par] = nullptr;
		}
		cudaGraph_t g = nullptr;
		uint64_t k = 0;
		const cudaStream_t keep = c->stream;
		c->stream = on; // the enqueue helpers launch on c->stream
		cudaError_t e = cudaStreamBeginCapture(on, cudaStreamCaptureModeThreadLocal);
		int rc = 1;
		if (e == cudaSuccess) {
			rc = enqueue_frame_kernels(c, true, false, &k, phase_mask);
			e = cudaStreamEndCapture(on, &g);
		}
		c->stream = keep;
		if (rc)
			return 1;
		if (e != cudaSuccess)
			return fail("frame phase capture: %s", cudaGetErrorString(e));
		CK(cudaGraphInstantiate(&c->phase_exec[phase_idx][par], g, 0));
		++c->rebuilds;
		cudaGraphDestroy(g);
		c->phase_kernels[phase_idx][par] = k;
		c->phase_valid[phase_idx][par] = true;
	}
	if (c->phase_kernels[phase_idx][par])
		CK(cudaGraphLaunch(c->phase_exec[phase_idx][par], on));
	c->launches += c->phase_kernels[phase_idx][par];
	return 0;
}

static int launch_frame_async(hgb_ctx *c) {
	const int par = c->cur;
	if (c->post_pending[par]) { // the tail of the frame before last compacts the visibility bytes this frame's levels are about to overwrite
		CK(cudaStreamWaitEvent(c->stream, c->ev_post_done[par], 0));
		c->post_pending[par] = false;
	}
	if (launch_phase_graph(c, 0, kPhasePre | kPhaseLevels, c->stream))
		return 1;
	CK(cudaEventRecord(c->ev_lvl_done, c->stream));
	CK(cudaStreamWaitEvent(c->stream2, c->ev_lvl_done, 0));
	if (launch_phase_graph(c, 2, kPhasePost, c->stream2)) // stream2 is in order: tails never overtake each other
		return 1;
	CK(cudaEventRecord(c->ev_post_done[par], c->stream2));
	c->post_pending[par] = true;
	return 0;
}

// graph path: one executable graph per (fused flag, world-buffer parity); re-captured after any edit (invalidate_graph)
static int launch_frame_graph(hgb_ctx *c, bool fused) {
	const int f = fused ? 1 : 0, par = c->cur;
	if (!c->graph_valid[f][par]) {
		if (c->graph_exec[f][par]) {
			cudaGraphExecDestroy(c->graph_exec[f][par]);
			c->graph_exec[f][par] = nullptr;
		}
		cudaGraph_t g = nullptr;
		uint64_t k = 0;
		CK(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
		const int rc = enqueue_frame_kernels(c, fused, false, &k);
		const cudaError_t e = cudaStreamEndCapture(c->stream, &g);
		if (rc)
			return 1;
		if (e != cudaSuccess)
			return fail("cudaStreamEndCapture: %s", cudaGetErrorString(e));
		CK(cudaGraphInstantiate(&c->graph_exec[f][par], g, 0));
		++c->rebuilds;
		cudaGraphDestroy(g);
		c->graph_kernels[f][par] = k;
		c->graph_valid[f][par] = true;
	}
	CK(cudaGraphLaunch(c->graph_exec[f][par], c->stream));
	c->launches += c->graph_kernels[f][par];
	return 0;
}

extern "C" {

int hgb_update(hgb_ctx *c, uint32_t n_play, const int64_t *play_t, int fused_cull) {
	if (!c)
		return fail("null context");
	if (!c->n_slots)
		return 0;
	CK(cudaSetDevice(c->device));
	if (fused_cull && !c->instance_mode && c->views.n_views == 0 && (c->ent[0].n_dev || c->ent[1].n_dev))
		return fail("hgb_update(fused_cull): no views set (hgb_set_views / hgb_set_instance_views)");
	if (prepare_tables(c))
		return 1;
	if (n_play > c->n_play)
		return fail("hgb_update: %u play clocks given but hgb_set_tracks declared %u play slots", n_play, c->n_play);

	// a2 StorePreviousWorldMatrices: pointer swap. a3 ReadyWorldMatrices: the 'updated' flags only exist as the preset mask.
	c->cur ^= 1;
	const bool with_preset = c->preset_pending;
	if (c->preset_dirty) {
		CK(cudaMemsetAsync(c->preset.p, 0, c->n_slots, c->stream));
		c->preset_dirty = false;
	}
	if (with_preset) {
		k_apply_presets<<<grid_for(c->n_preset, 128), 128, 0, c->stream>>>(c->preset_pos.as<uint32_t>(), c->preset_mat.as<float4>(), c->n_preset,
			c->world[c->cur].as<float4>(), c->preset.as<uint8_t>());
		++c->launches;
		c->preset_pending = false;
		c->preset_dirty = true;
	}
	if (c->n_play) {
		// the clocks go up from a ring of pinned slots (a plain DMA, nothing staged by the runtime); a slot is reused only once the copy
		// that read it has run (its event: long done in practice, the ring is 8 frames deep)
		const uint32_t slot = c->clock_slot++ % kClockRing;
		if (c->clock_used[slot])
			CK(cudaEventSynchronize(c->clock_ev[slot]));
		int64_t *dst = c->play_t_h.as<int64_t>() + size_t(slot) * c->n_play;
		for (uint32_t i = 0; i < c->n_play; ++i)
			dst[i] = i < n_play ? play_t[i] : INT64_MIN;
		CK(cudaMemcpyAsync(c->play_t_d.p, dst, size_t(c->n_play) * 8, cudaMemcpyHostToDevice, c->stream));
		CK(cudaEventRecord(c->clock_ev[slot], c->stream));
		c->clock_used[slot] = true;
	}

	if (!c->en_target_slot_h.empty() && c->n_play)
		c->en_state_dirty = true;
	const bool fused = fused_cull != 0;
	if (c->async_auto)
		c->sync_streak = c->result_read ? std::min(c->sync_streak + 1u, 1000u) : 0u;
	c->result_read = false;
	// (with the peer gather only for batches of at least 4 M nodes: on a short frame the gather's fixed latencies are longer than the
	// frame and the tail is better off the frame's stream — measured: c4, 16 M nodes per GPU, e2e + 7 / + 9 / + 25 % at N = 2 / 4 / 8 with
	// the one-graph frame; c5 strong + 7 % at 8.4 M nodes per GPU, equal at 4.2 M, - 25 % at 2.1 M; profiles/README.md r02)
	const bool sync_pays = !c->peer.on || c->n_pos >= (4u << 20);
	c->last_frame_async = c->async_post && !(c->async_auto && c->sync_streak >= 2 && sync_pays) && c->use_graph && fused && !with_preset && !c->prof.on &&
						  (c->ent[0].n_dev || c->ent[1].n_dev || c->peer.on);
	if (c->last_frame_async)
		return launch_frame_async(c);
	join_post(c);
	if (!c->use_graph || with_preset || c->prof.on) {
		uint64_t k = 0;
		if (enqueue_frame_kernels(c, fused, with_preset, &k))
			return 1;
		c->launches += k;
		return 0;
	}
	return launch_frame_graph(c, fused);
}

int hgb_cull(hgb_ctx *c) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaSetDevice(c->device));
	if (!c->instance_mode && c->views.n_views == 0)
		return fail("hgb_cull: no views set");
	if (prepare_tables(c))
		return 1;
	uint64_t k = 0;
	for (int kind = 0; kind < 2; ++kind) {
		EntryTable &t = c->ent[kind];
		if (!t.n)
			continue;
		CullDev d{};
		d.mtx_pos = t.mtx_pos.as<uint32_t>();
		d.meta = t.meta.as<uint32_t>();
		d.view = t.view_d.as<uint16_t>();
		d.world = c->world[c->cur].as<float>();
		d.bounds6 = c->bounds.as<float>();
		d.inst_planes = c->inst_planes.as<float4>();
		d.vis = t.vis[c->cur].as<uint8_t>();
		d.n = t.n;
		prof_begin(c, HGB_K_CULL_FLAGS);
		if (c->instance_mode)
			k_cull_flags<true><<<grid_for(t.n), 256, 0, c->stream>>>(d, c->views);
		else
			k_cull_flags<false><<<grid_for(t.n), 256, 0, c->stream>>>(d, c->views);
		prof_end(c);
		++k;
	}
	CK(cudaGetLastError());
	if (enqueue_compaction(c, &k))
		return 1;
	c->launches += k;
	return 0;
}

int hgb_cull_list(hgb_ctx *c, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, uint32_t n_bounds, const float *bounds6, const float *frustum24,
	uint32_t *out_idx, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (out_count)
		*out_count = 0;
	if (!n)
		return 0;
	if (n_bounds > 0x00ffffffu) // the entry meta word keeps the bounds index in its low 24 bits, flags above (like hgb_set_bounds)
		return fail("hgb_cull_list: %u bounds exceed the 2^24-1 the entry word can index", n_bounds);
	CK(cudaSetDevice(c->device));
	std::vector<uint32_t> pos(n);
	for (uint32_t e = 0; e < n; ++e) {
		if (mtx_slot[e] >= c->n_slots)
			return fail("hgb_cull_list: entry %u references slot %u outside the topology", e, mtx_slot[e]);
		if (bounds_idx[e] >= n_bounds)
			return fail("hgb_cull_list: entry %u references bounds %u outside the table", e, bounds_idx[e]);
		pos[e] = c->pos_of_slot[mtx_slot[e]];
	}
	const uint32_t n_tiles = (n + kCompactTile - 1) / kCompactTile;
	float cond[24];
	precondition_planes(frustum24, 6, cond);
	interleave_bounds(bounds6, n_bounds, c->bounds_stage); // stays alive until the stream synchronisation below
	if (upload(c, c->adhoc_pos, pos.data(), size_t(n) * 4) || upload(c, c->adhoc_meta, bounds_idx, size_t(n) * 4) ||
		upload(c, c->adhoc_bounds, c->bounds_stage.data(), size_t(n_bounds) * 24) || upload(c, c->adhoc_planes, cond, 96))
		return 1;
	CK(c->adhoc_vis.reserve(size_t(n_tiles + 1) * kCompactTile));
	CK(c->adhoc_out.reserve(size_t(n) * 4));
	CK(c->adhoc_scratch.reserve(size_t(n_tiles + 2) * 4));
	CK(cudaMemsetAsync(c->adhoc_vis.p, 0, size_t(n_tiles + 1) * kCompactTile, c->stream));
	CK(cudaMemsetAsync(c->adhoc_scratch.p, 0, size_t(n_tiles + 2) * 4, c->stream));
	CullDev d{};
	d.mtx_pos = c->adhoc_pos.as<uint32_t>();
	d.meta = c->adhoc_meta.as<uint32_t>();
	d.world = c->world[c->cur].as<float>();
	d.bounds6 = c->adhoc_bounds.as<float>();
	d.vis = c->adhoc_vis.as<uint8_t>();
	d.n = n;
	Views v{c->adhoc_planes.as<float4>(), 1};
	prof_begin(c, HGB_K_CULL_FLAGS);
	k_cull_flags<false><<<grid_for(n), 256, 0, c->stream>>>(d, v);
	prof_end(c);
	CompactDev cd{};
	cd.vis = d.vis;
	cd.n = n;
	cd.bit = 0;
	cd.out = c->adhoc_out.as<uint32_t>();
	cd.count = c->adhoc_scratch.as<uint32_t>();
	cd.ticket = c->adhoc_scratch.as<uint32_t>() + 1;
	cd.status = c->adhoc_scratch.as<uint32_t>() + 2;
	prof_begin(c, HGB_K_COMPACT);
	k_compact<<<n_tiles, kCompactThreads, 0, c->stream>>>(cd);
	prof_end(c);
	c->launches += 2;
	CK(cudaGetLastError());
	uint32_t count = 0;
	CK(cudaMemcpyAsync(&count, cd.count, 4, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream)); // also keeps pos[] alive until the upload is done
	if (count > n)
		return fail("hgb_cull_list: internal error (count %u > n %u)", count, n);
	if (count)
		CK(cudaMemcpyAsync(out_idx, c->adhoc_out.p, size_t(count) * 4, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	if (out_count)
		*out_count = count;
	return 0;
}

// Opt-in culling of skinned display lists (see k_cull_skinned): stateless like hgb_cull_list. List i owns the bones
// [bone_off[i], bone_off[i+1]) = (mtx_slot, bind_idx) pairs in SkinnedModelDisplayList order; bounds_idx[i] selects the list's MinMax.
int hgb_cull_skinned_list(hgb_ctx *c, uint32_t n_lists, const uint32_t *bone_off, const uint32_t *mtx_slot, const uint32_t *bind_idx, uint32_t n_bind,
	const float *bind_pose12, const uint32_t *bounds_idx, uint32_t n_bounds, const float *bounds6, const float *frustum24, uint32_t *out_idx, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (out_count)
		*out_count = 0;
	if (!n_lists)
		return 0;
	CK(cudaSetDevice(c->device));
	const uint32_t n_bones = bone_off[n_lists];
	std::vector<uint32_t> pos(std::max(n_bones, 1u));
	for (uint32_t i = 0; i < n_lists; ++i) {
		if (bone_off[i + 1] < bone_off[i] || bone_off[i + 1] - bone_off[i] > 32)
			return fail("hgb_cull_skinned_list: list %u has %u bones (at most 32, render_pipeline.h:38)", i, bone_off[i + 1] - bone_off[i]);
		if (bounds_idx[i] >= n_bounds)
			return fail("hgb_cull_skinned_list: list %u references bounds %u outside the table", i, bounds_idx[i]);
	}
	for (uint32_t b = 0; b < n_bones; ++b) {
		if (mtx_slot[b] >= c->n_slots)
			return fail("hgb_cull_skinned_list: bone %u references slot %u outside the topology", b, mtx_slot[b]);
		if (bind_idx[b] >= n_bind)
			return fail("hgb_cull_skinned_list: bone %u references bind pose %u outside the pool", b, bind_idx[b]);
		pos[b] = c->pos_of_slot[mtx_slot[b]];
	}
	float cond[24];
	precondition_planes(frustum24, 6, cond);
	interleave_bounds(bounds6, n_bounds, c->bounds_stage);
	DBuf d_off, d_pos, d_bind, d_pose, d_bidx, d_vis;
	int rc = upload(c, d_off, bone_off, size_t(n_lists + 1) * 4) || upload(c, d_pos, pos.data(), size_t(n_bones) * 4) || upload(c, d_bind, bind_idx, size_t(n_bones) * 4) ||
			 upload(c, d_pose, bind_pose12, size_t(n_bind) * 48) || upload(c, d_bidx, bounds_idx, size_t(n_lists) * 4) ||
			 upload(c, c->adhoc_bounds, c->bounds_stage.data(), size_t(n_bounds) * 24) || upload(c, c->adhoc_planes, cond, 96);
	if (!rc && d_vis.reserve(n_lists) != cudaSuccess)
		rc = fail("hgb_cull_skinned_list: out of device memory");
	std::vector<uint8_t> vis(n_lists);
	if (!rc) {
		SkinCullDev d{};
		d.bone_off = d_off.as<uint32_t>(), d.mtx_pos = d_pos.as<uint32_t>(), d.bind_idx = d_bind.as<uint32_t>(), d.bind_pose = d_pose.as<float>();
		d.bounds_idx = d_bidx.as<uint32_t>(), d.bounds6 = c->adhoc_bounds.as<float>(), d.world = c->world[c->cur].as<float>();
		d.planes = c->adhoc_planes.as<float4>(), d.vis = d_vis.as<uint8_t>(), d.n_lists = n_lists;
		k_cull_skinned<<<grid_for(n_lists * 32), 256, 0, c->stream>>>(d);
		++c->launches;
		cudaError_t e = cudaGetLastError();
		if (e == cudaSuccess)
			e = cudaMemcpyAsync(vis.data(), d_vis.p, n_lists, cudaMemcpyDeviceToHost, c->stream);
		if (e == cudaSuccess)
			e = cudaStreamSynchronize(c->stream);
		if (e != cudaSuccess)
			rc = fail("hgb_cull_skinned_list: %s", cudaGetErrorString(e));
	} else {
		cudaStreamSynchronize(c->stream);
	}
	DBuf *tmp[] = {&d_off, &d_pos, &d_bind, &d_pose, &d_bidx, &d_vis};
	for (DBuf *b : tmp)
		b->release();
	if (rc)
		return 1;
	uint32_t n = 0;
	for (uint32_t i = 0; i < n_lists; ++i)
		if (vis[i])
			out_idx[n++] = i; // order-preserving, like std::remove_if
	if (out_count)
		*out_count = n;
	return 0;
}

int hgb_bone_matrices(hgb_ctx *c, int with_previous) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (!c->n_bones)
		return 0;
	CK(cudaSetDevice(c->device));
	if (prepare_tables(c))
		return 1;
	BoneDev b{};
	b.mtx_pos = c->bone_mtx_pos.as<uint32_t>();
	b.bind_idx = c->bone_bind_idx.as<uint32_t>();
	b.bind_pose = c->bind_pose.as<float>();
	b.world = c->world[c->cur].as<float>();
	b.world_prev = with_previous ? c->world[c->cur ^ 1].as<float>() : nullptr;
	b.out = c->bone_out[0].as<float>();
	b.out_prev = c->bone_out[1].as<float>();
	b.n = c->n_bones;
	prof_begin(c, HGB_K_BONE_MATRICES);
	k_bone_matrices<<<grid_for(b.n), 256, 0, c->stream>>>(b);
	prof_end(c);
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

// Scene::GetMinMax / GetNodesMinMax: objects are resolved to positions and sorted by position (the union is order-independent), so the
// world rows are read in ascending address order
static int resolve_objects(hgb_ctx *c) {
	if (!c->obj_dirty)
		return 0;
	const uint32_t n = c->n_obj;
	std::vector<uint64_t> order(n);
	for (uint32_t i = 0; i < n; ++i) {
		const uint32_t sl = c->obj_slot_h[i];
		uint32_t p = kRoot;
		if (sl != HGB_ROOT) {
			if (sl >= c->n_slots || !pos_is_live(c, c->pos_of_slot[sl]))
				return fail("hgb_minmax: object %u references slot %u which is not a live transform", i, sl);
			p = c->pos_of_slot[sl];
		}
		order[i] = (uint64_t(p) << 32) | i;
	}
	std::sort(order.begin(), order.end());
	std::vector<uint32_t> pos(n), meta(n), seg(c->obj_seg_h.empty() ? 0 : n);
	for (uint32_t k = 0; k < n; ++k) {
		const uint32_t i = uint32_t(order[k]);
		pos[k] = uint32_t(order[k] >> 32);
		meta[k] = c->obj_ob_h[i];
		if (!seg.empty())
			seg[k] = c->obj_seg_h[i];
	}
	if (upload(c, c->obj_pos, pos.data(), size_t(n) * 4) || upload(c, c->obj_meta, meta.data(), size_t(n) * 4) ||
		upload(c, c->obj_seg, seg.data(), seg.size() * 4))
		return 1;
	CK(c->mm_acc.reserve(std::max<size_t>(c->n_seg, 1) * 32));
	CK(cudaStreamSynchronize(c->stream));
	c->obj_dirty = false;
	return 0;
}

int hgb_minmax(hgb_ctx *c) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaSetDevice(c->device));
	if (resolve_objects(c))
		return 1;
	if (!c->n_seg)
		return 0;
	k_minmax_init<<<grid_for(c->n_seg * 8), 256, 0, c->stream>>>(c->mm_acc.as<uint32_t>(), c->n_seg);
	++c->launches;
	if (c->n_obj) {
		MinMaxDev d{};
		d.obj_pos = c->obj_pos.as<uint32_t>();
		d.obj_meta = c->obj_meta.as<uint32_t>();
		d.obj_seg = c->obj_seg_h.empty() ? nullptr : c->obj_seg.as<uint32_t>();
		d.world = c->world[c->cur].as<float>();
		d.obounds6 = c->obounds.as<float>();
		d.acc = c->mm_acc.as<uint32_t>();
		d.n = c->n_obj;
		// persistent grid: 8 CTAs of 8 warps per SM at most, every warp a contiguous chunk (multiple of 32 objects)
		const uint32_t max_warps = 148u * 8u * 8u;
		uint32_t per_warp = (d.n + max_warps - 1) / max_warps;
		per_warp = std::max(32u, (per_warp + 31u) & ~31u);
		const uint32_t n_warps = (d.n + per_warp - 1) / per_warp;
		prof_begin(c, HGB_K_MINMAX);
		k_minmax<<<(n_warps + 7) / 8, 256, 0, c->stream>>>(d, per_warp);
		prof_end(c);
		++c->launches;
	}
	CK(cudaGetLastError());
	return 0;
}

int hgb_get_minmax(hgb_ctx *c, float *out6, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (!c->n_seg)
		return 0;
	if (!c->mm_acc.p || c->obj_dirty)
		return fail("hgb_get_minmax: hgb_minmax has not run since the objects or the hierarchy changed");
	CK(cudaSetDevice(c->device));
	CK(c->staging.reserve(size_t(c->n_seg) * 32));
	CK(cudaMemcpyAsync(c->staging.p, c->mm_acc.p, size_t(c->n_seg) * 32, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	const uint32_t *a = c->staging.as<uint32_t>();
	for (uint32_t s = 0; s < c->n_seg; ++s) {
		for (int j = 0; j < 6; ++j)
			out6[size_t(s) * 6 + j] = key_float(a[size_t(s) * 8 + j]);
		if (out_count)
			out_count[s] = a[size_t(s) * 8 + 6];
	}
	return 0;
}

// ComputeModelDisplayListSortKeys for the visible list (kind, view) of the last cull
int hgb_sort_keys(hgb_ctx *c, int kind, uint32_t view, const float *view_mat12) {
	if (!c)
		return fail("null context");
	join_post(c);
	if ((kind != 0 && kind != 1) || view >= uint32_t(kMaxViews) || !view_mat12)
		return fail("hgb_sort_keys: bad kind/view/matrix");
	EntryTable &t = c->ent[kind];
	if (!t.n)
		return 0;
	CK(cudaSetDevice(c->device));
	if (prepare_tables(c))
		return 1;
	if (!t.out[view].p)
		return fail("hgb_sort_keys: list (kind %d, view %u) has not been culled", kind, view);
	CK(t.keys[view].reserve(size_t(t.n) * 4));
	SortKeyDev d{};
	d.visible = t.out[view].as<uint32_t>();
	d.count = t.counts.as<uint32_t>() + view;
	d.mtx_pos = t.mtx_pos.as<uint32_t>();
	d.meta = t.meta.as<uint32_t>();
	d.world = c->world[c->cur].as<float>();
	d.bounds6 = c->bounds.as<float>();
	memcpy(d.view, view_mat12, 48);
	d.keys = t.keys[view].as<uint32_t>();
	prof_begin(c, HGB_K_SORT_KEYS);
	k_sort_keys<<<grid_for(t.n), 256, 0, c->stream>>>(d, t.n);
	prof_end(c);
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

int hgb_get_sort_keys(hgb_ctx *c, int kind, uint32_t view, uint32_t *out_keys, uint32_t capacity, uint32_t *out_count) {
	uint32_t n = 0;
	if (hgb_get_visible_count(c, kind, view, &n))
		return 1;
	if (out_count)
		*out_count = n;
	if (n > capacity)
		return fail("hgb_get_sort_keys: %u keys but capacity %u", n, capacity);
	if (n) {
		if (!c->ent[kind].keys[view].p)
			return fail("hgb_get_sort_keys: hgb_sort_keys has not run for (kind %d, view %u)", kind, view);
		CK(cudaMemcpyAsync(out_keys, c->ent[kind].keys[view].p, size_t(n) * 4, cudaMemcpyDeviceToHost, c->stream));
		CK(cudaStreamSynchronize(c->stream));
	}
	return 0;
}

// ---- getters -------------------------------------------------------------------------------------------------------------
// slots the device arrays hold: the topology's slot count plus the append slack of the last depth sort (hgb_append_nodes takes free slots below it)
int hgb_slot_capacity(hgb_ctx *c, uint32_t *out_capacity) {
	if (!c)
		return fail("null context");
	*out_capacity = slot_limit(c);
	return 0;
}

int hgb_get_worlds(hgb_ctx *c, uint32_t first, uint32_t count, float *out12, int previous) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (uint64_t(first) + count > slot_limit(c))
		return fail("hgb_get_worlds: range outside the topology");
	if (!count)
		return 0;
	CK(cudaSetDevice(c->device));
	const float *w = c->world[previous ? c->cur ^ 1 : c->cur].as<float>();
	if (c->identity) {
		CK(cudaMemcpyAsync(out12, w + size_t(first) * 12, size_t(count) * 48, cudaMemcpyDeviceToHost, c->stream));
	} else {
		CK(c->staging_d.reserve(size_t(count) * 48));
		k_gather_worlds<<<grid_for(count * 3), 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(w), c->pos_of_slot_d.as<uint32_t>(), first, count,
			c->staging_d.as<float4>());
		++c->launches;
		CK(cudaMemcpyAsync(out12, c->staging_d.p, size_t(count) * 48, cudaMemcpyDeviceToHost, c->stream));
	}
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

int hgb_get_trs(hgb_ctx *c, uint32_t first, uint32_t count, float *out9) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (uint64_t(first) + count > slot_limit(c))
		return fail("hgb_get_trs: range outside the topology");
	if (!count)
		return 0;
	CK(cudaSetDevice(c->device));
	CK(c->staging_d.reserve(size_t(count) * 36));
	k_gather_trs<<<grid_for(count), 256, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), first, count,
		c->staging_d.as<float>(), c->anim_val.as<float4>());
	++c->launches;
	CK(cudaMemcpyAsync(out9, c->staging_d.p, size_t(count) * 36, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

int hgb_get_visible_count(hgb_ctx *c, int kind, uint32_t view, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	c->result_read = true;
	join_post(c);
	if ((kind != 0 && kind != 1) || view >= uint32_t(kMaxViews))
		return fail("hgb_get_visible_count: bad kind/view");
	*out_count = 0;
	if (!c->ent[kind].n)
		return 0;
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(out_count, c->ent[kind].counts.as<uint32_t>() + view, 4, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

// The list comes down in ONE round trip: its length is only known on the device, so the copy is sized from the previous call's count
// (+25 %, at least 4096 entries, at most the capacity) and goes out together with the 4-byte count, one stream synchronisation for
// both; only when the list outgrew the guess does a second copy fetch the rest.
static int fetch_visible(hgb_ctx *c, int kind, uint32_t view, uint32_t *out_idx, uint32_t capacity, uint32_t *out_count) {
	if ((kind != 0 && kind != 1) || view >= uint32_t(kMaxViews))
		return fail("hgb_get_visible: bad kind/view");
	EntryTable &t = c->ent[kind];
	if (out_count)
		*out_count = 0;
	if (!t.n)
		return 0;
	CK(cudaSetDevice(c->device));
	CK(c->count_h.reserve(64));
	uint32_t *cnt = c->count_h.as<uint32_t>();
	const uint64_t want = uint64_t(t.last_count[view]) + t.last_count[view] / 16 + 4096; // a frame rarely moves the count by more than a few per cent
	const uint32_t guess = uint32_t(std::min<uint64_t>(std::min<uint64_t>(want, capacity), t.n));
	CK(cudaMemcpyAsync(cnt, t.counts.as<uint32_t>() + view, 4, cudaMemcpyDeviceToHost, c->stream));
	if (guess)
		CK(cudaMemcpyAsync(out_idx, t.out[view].p, size_t(guess) * 4, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	const uint32_t n = *cnt;
	if (out_count)
		*out_count = n;
	if (n > capacity)
		return fail("hgb_get_visible: %u survivors but capacity %u", n, capacity);
	if (n > guess) {
		CK(cudaMemcpyAsync(out_idx + guess, t.out[view].as<uint32_t>() + guess, size_t(n - guess) * 4, cudaMemcpyDeviceToHost, c->stream));
		CK(cudaStreamSynchronize(c->stream));
	}
	t.last_count[view] = n;
	return 0;
}

int hgb_get_visible(hgb_ctx *c, int kind, uint32_t view, uint32_t *out_idx, uint32_t capacity, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	c->result_read = true;
	join_post(c);
	return fetch_visible(c, kind, view, out_idx, capacity, out_count);
}

int hgb_get_visible_ptr(hgb_ctx *c, int kind, uint32_t view, const uint32_t **out_idx, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	c->result_read = true;
	join_post(c);
	if ((kind != 0 && kind != 1) || view >= uint32_t(kMaxViews))
		return fail("hgb_get_visible_ptr: bad kind/view");
	CK(cudaSetDevice(c->device));
	HBuf &h = c->visible_h[kind];
	CK(h.reserve(std::max<size_t>(c->ent[kind].n, 1) * 4)); // pinned, owned by the context: the copy is a plain DMA
	*out_idx = h.as<uint32_t>();
	return fetch_visible(c, kind, view, h.as<uint32_t>(), c->ent[kind].n, out_count);
}

// One opaque 16-byte record per display entry of a kind (the facade stores its ModelDisplayList there: material pointer, mtx_idx,
// mdl_idx, lst_idx — render_pipeline.h:577-582), so that the visible list can come back as the records themselves. Records
// [first, first + count): the whole table after hgb_set_display_entries, the new tail after hgb_append_display_entries.
int hgb_set_entry_records(hgb_ctx *c, int kind, uint32_t first, uint32_t count, const void *rec16) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (kind != 0 && kind != 1)
		return fail("hgb_set_entry_records: bad kind");
	EntryTable &t = c->ent[kind];
	if (uint64_t(first) + count > t.n)
		return fail("hgb_set_entry_records: records [%u, %u) for a table of %u entries (hgb_set_display_entries first)", first, first + count, t.n);
	if (first > t.n_records)
		return fail("hgb_set_entry_records: records [%u, %u) leave a gap behind the %u already set", first, first + count, t.n_records);
	if (!count)
		return 0;
	CK(cudaSetDevice(c->device));
	const size_t cap = size_t(std::max(t.n_dev, t.n)) * 16;
	if (t.records.cap < cap) { // grows only with the table (hgb_set_display_entries resets n_records)
		if (first)
			return fail("hgb_set_entry_records: the table grew past its slack; set all records again");
		CK(t.records.reserve(cap));
	}
	CK(cudaMemcpyAsync(t.records.as<uint8_t>() + size_t(first) * 16, rec16, size_t(count) * 16, cudaMemcpyHostToDevice, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	t.n_records = std::max(t.n_records, first + count);
	return 0;
}

// Pipelined form: the records land in chunks of `chunk_records`, each with its own event, so the caller copies chunk i out of the pinned
// buffer while chunk i + 1 is still on the link. Returns once the COUNT is known (one short wait behind the frame's tail); chunk i is
// complete when hgb_visible_records_wait(ctx, kind, i) returns (callable from any thread). The size guess follows the last count; a
// list that outgrew it gets its remaining chunks issued before this returns.
int hgb_get_visible_records_begin(hgb_ctx *c, int kind, uint32_t view, uint32_t chunk_records, const void **out_rec16, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	c->result_read = true;
	join_post(c);
	if ((kind != 0 && kind != 1) || view >= uint32_t(kMaxViews))
		return fail("hgb_get_visible_records: bad kind/view");
	if (chunk_records < 1024)
		return fail("hgb_get_visible_records: chunks of at least 1024 records");
	EntryTable &t = c->ent[kind];
	*out_count = 0;
	*out_rec16 = nullptr;
	t.rec_chunks = 0;
	if (!t.n)
		return 0;
	if (t.n_records != t.n || !t.records.p)
		return fail("hgb_get_visible_records: no records for this table (hgb_set_entry_records)");
	CK(cudaSetDevice(c->device));
	const uint32_t max_chunks = (t.n + chunk_records - 1) / chunk_records;
	CK(t.records_out.reserve(size_t(max_chunks) * chunk_records * 16)); // whole chunks travel: the tail of the last one is scratch
	CK(t.records_h.reserve(size_t(max_chunks) * chunk_records * 16));
	CK(c->count_h.reserve(64));
	if (!t.rec_count_ev)
		CK(cudaEventCreateWithFlags(&t.rec_count_ev, cudaEventDisableTiming));
	uint32_t *cnt = c->count_h.as<uint32_t>();
	k_gather_records<<<148 * 8, 256, 0, c->stream>>>(t.records.as<uint4>(), t.out[view].as<uint32_t>(), t.counts.as<uint32_t>() + view, t.n, t.records_out.as<uint4>());
	++c->launches;
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(cnt, t.counts.as<uint32_t>() + view, 4, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaEventRecord(t.rec_count_ev, c->stream));
	auto issue = [&](uint32_t upto) -> int { // chunks [t.rec_chunks, upto)
		while (t.rec_ev.size() < upto) {
			cudaEvent_t e = nullptr;
			CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
			t.rec_ev.push_back(e);
		}
		for (uint32_t i = t.rec_chunks; i < upto; ++i) {
			const size_t off = size_t(i) * chunk_records * 16;
			CK(cudaMemcpyAsync(static_cast<uint8_t *>(t.records_h.p) + off, t.records_out.as<uint8_t>() + off, size_t(chunk_records) * 16, cudaMemcpyDeviceToHost, c->stream));
			CK(cudaEventRecord(t.rec_ev[i], c->stream));
		}
		t.rec_chunks = std::max(t.rec_chunks, upto);
		return 0;
	};
	const uint64_t want = uint64_t(t.last_count[view]) + t.last_count[view] / 32 + 1024; // 16 bytes a record: a tighter guess than the index fetch
	const uint32_t guess_chunks = uint32_t(std::min<uint64_t>((want + chunk_records - 1) / chunk_records, max_chunks));
	if (issue(guess_chunks))
		return 1;
	CK(cudaEventSynchronize(t.rec_count_ev));
	const uint32_t n = *cnt;
	if (n > t.n)
		return fail("hgb_get_visible_records: internal error (count %u > entries %u)", n, t.n);
	if (issue((n + chunk_records - 1) / chunk_records))
		return 1;
	t.last_count[view] = n;
	*out_count = n;
	*out_rec16 = t.records_h.p;
	return 0;
}

int hgb_visible_records_wait(hgb_ctx *c, int kind, uint32_t chunk) {
	if (!c || (kind != 0 && kind != 1) || chunk >= c->ent[kind].rec_chunks)
		return 1; // (no message: may be called from worker threads)
	return cudaEventSynchronize(c->ent[kind].rec_ev[chunk]) == cudaSuccess ? 0 : 1;
}

// The visible list of (kind, view) as the 16-byte records of the surviving entries, in order, in a pinned buffer owned by the context
// (valid until the next call for this kind). Gathered on the device, one round trip like hgb_get_visible_ptr.
int hgb_get_visible_records_ptr(hgb_ctx *c, int kind, uint32_t view, const void **out_rec16, uint32_t *out_count) {
	if (hgb_get_visible_records_begin(c, kind, view, 1u << 16, out_rec16, out_count))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

int hgb_get_instance_offsets(hgb_ctx *c, uint32_t *out) {
	if (!c)
		return fail("null context");
	c->result_read = true;
	join_post(c);
	if (!c->instance_mode)
		return fail("hgb_get_instance_offsets: not in instance mode");
	CK(cudaSetDevice(c->device));
	if (c->n_inst)
		CK(cudaMemcpyAsync(out, c->inst_rank.p, size_t(c->n_inst) * 4, cudaMemcpyDeviceToHost, c->stream));
	uint32_t total = 0;
	if (hgb_get_visible_count(c, 0, 0, &total))
		return 1;
	out[c->n_inst] = total;
	return 0;
}

int hgb_get_bone_matrices(hgb_ctx *c, float *out16, int previous) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (!c->n_bones)
		return 0;
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(out16, c->bone_out[previous ? 1 : 0].p, size_t(c->n_bones) * 64, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

void *hgb_device_ptr(hgb_ctx *c, int which) {
	if (!c)
		return nullptr;
	switch (which) {
	case 0: return c->world[c->cur].p;
	case 1: return c->world[c->cur ^ 1].p;
	case 2: return c->ent[0].out[0].p;
	case 3: return c->ent[0].counts.p;
	case 4: return c->bone_out[0].p;
	case 5: return c->inst_rank.p;
	case 6: return c->tiles.p;
	case 7: return c->peer_out_list.p;
	case 8: return c->peer_rank_off.p;
	default: return nullptr;
	}
}

} // extern "C"
