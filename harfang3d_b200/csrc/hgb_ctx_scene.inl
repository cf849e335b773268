// Edit-time half of the C ABI: hierarchy -> depth-sorted layout, uploads, display-entry tables, skins, animation pools, views.
// Included by hgb_ctx.cu.

static int upload(hgb_ctx *c, DBuf &dst, const void *src, size_t bytes) {
	CK(dst.reserve(bytes ? bytes : 4));
	if (bytes)
		CK(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
	return 0;
}

// plane {x,y,z,w} -> {x,y,z,-2w}: TestVisibility compares against plane_d_x2 = -plane.w * 2.f (foundation/frustum.cpp:72), exact in float
static void precondition_planes(const float *in, uint32_t n_planes, float *out) {
	for (uint32_t i = 0; i < n_planes; ++i) {
		out[i * 4 + 0] = in[i * 4 + 0];
		out[i * 4 + 1] = in[i * 4 + 1];
		out[i * 4 + 2] = in[i * 4 + 2];
		out[i * 4 + 3] = -in[i * 4 + 3] * 2.f;
	}
}

static inline dim3 grid_for(uint32_t n, uint32_t block = 256) { return dim3((n + block - 1) / block); }

extern "C" {

int hgb_set_topology(hgb_ctx *c, uint32_t n, const uint32_t *parent_slot, const uint8_t *used) {
	if (!c)
		return fail("null context");
	if (n >= 0x00ffffffu)
		return fail("hgb_set_topology: %u slots exceeds the reference's vector_list capacity (2^24-1, foundation/vector_list.h:40)", n);
	CK(cudaSetDevice(c->device));

	// children lists (CSR by parent slot). A parent that is unused/out of range makes the child a root (scene.cpp:268-272).
	std::vector<uint32_t> child_cnt(n + 1, 0), roots;
	auto is_used = [&](uint32_t s) { return !used || used[s]; };
	auto parent_of = [&](uint32_t s) -> uint32_t {
		const uint32_t p = parent_slot ? parent_slot[s] : HGB_ROOT;
		return (p < n && is_used(p) && p != s) ? p : HGB_ROOT;
	};
	uint32_t n_used = 0;
	for (uint32_t s = 0; s < n; ++s) {
		if (!is_used(s))
			continue;
		++n_used;
		const uint32_t p = parent_of(s);
		if (p == HGB_ROOT)
			roots.push_back(s);
		else
			++child_cnt[p + 1];
	}
	for (uint32_t s = 0; s < n; ++s)
		child_cnt[s + 1] += child_cnt[s];
	std::vector<uint32_t> child(n_used), fill(child_cnt.begin(), child_cnt.end() - 1);
	for (uint32_t s = 0; s < n; ++s)
		if (is_used(s)) {
			const uint32_t p = parent_of(s);
			if (p != HGB_ROOT)
				child[fill[p]++] = s;
		}

	// level-by-level order: roots in slot order, then the children of each node of the previous level, in order
	std::vector<uint32_t> slot_of_pos;
	slot_of_pos.reserve(n);
	std::vector<uint32_t> level_off{0};
	slot_of_pos.insert(slot_of_pos.end(), roots.begin(), roots.end());
	size_t lvl_begin = 0;
	while (lvl_begin < slot_of_pos.size()) {
		const size_t lvl_end = slot_of_pos.size();
		level_off.push_back(uint32_t(lvl_end));
		for (size_t i = lvl_begin; i < lvl_end; ++i) {
			const uint32_t s = slot_of_pos[i];
			for (uint32_t k = child_cnt[s]; k < child_cnt[s + 1]; ++k)
				slot_of_pos.push_back(child[k]);
		}
		lvl_begin = lvl_end;
	}
	if (slot_of_pos.size() != n_used)
		return fail("hgb_set_topology: parent cycle detected (%zu of %u transforms reachable from roots); the reference rejects cycles in "
					"SetTransformParent (component.cpp:184-193)",
			slot_of_pos.size(), n_used);
	const uint32_t n_pos = n_used;
	for (uint32_t s = 0; s < n; ++s)
		if (!is_used(s))
			slot_of_pos.push_back(s); // holes live after the last level and are never computed

	std::vector<uint32_t> pos_of_slot(n);
	bool identity = true;
	for (uint32_t p = 0; p < n; ++p) {
		pos_of_slot[slot_of_pos[p]] = p;
		identity = identity && slot_of_pos[p] == p;
	}
	std::vector<uint32_t> parent_pos(n, kRoot);
	for (uint32_t p = 0; p < n_pos; ++p) {
		const uint32_t par = parent_of(slot_of_pos[p]);
		parent_pos[p] = par == HGB_ROOT ? kRoot : pos_of_slot[par];
	}

	// carry the device TRS over to the new order (animation may have modified it on the device)
	const uint32_t keep = std::min(c->n_slots, n);
	const bool carry = keep > 0 && c->trs.p;
	if (carry) {
		CK(c->staging_d.reserve(size_t(keep) * 9 * 4));
		k_gather_trs<<<grid_for(keep), 256, 0, c->stream>>>(c->trs.as<float>(), c->trs_stride, c->identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), 0, keep,
			c->staging_d.as<float>());
		++c->launches;
		CK(cudaStreamSynchronize(c->stream));
	}

	const size_t stride = (size_t(n) + 31) & ~size_t(31);
	DBuf new_trs;
	CK(new_trs.reserve(std::max<size_t>(stride, 32) * 9 * 4));
	CK(cudaMemsetAsync(new_trs.p, 0, new_trs.cap, c->stream));
	if (upload(c, c->pos_of_slot_d, pos_of_slot.data(), size_t(n) * 4) || upload(c, c->parent_pos, parent_pos.data(), size_t(n) * 4))
		return 1;
	if (carry) {
		k_scatter_trs<<<grid_for(keep), 256, 0, c->stream>>>(c->staging_d.as<float>(), identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), 0, keep,
			new_trs.as<float>(), stride);
		++c->launches;
	}
	CK(cudaStreamSynchronize(c->stream)); // host vectors above go out of scope
	c->trs.release();
	c->trs = new_trs;
	c->trs_stride = stride;
	for (int i = 0; i < 2; ++i) {
		const size_t bytes = std::max<size_t>(n, 1) * 48;
		if (c->world[i].cap < bytes) {
			CK(c->world[i].reserve(bytes));
			CK(cudaMemsetAsync(c->world[i].p, 0, c->world[i].cap, c->stream));
		}
	}
	CK(c->preset.reserve(std::max<size_t>(n, 4)));
	CK(cudaMemsetAsync(c->preset.p, 0, c->preset.cap, c->stream));
	c->preset_dirty = false;

	c->n_slots = n;
	c->n_pos = n_pos;
	c->pos_of_slot.swap(pos_of_slot);
	c->slot_of_pos.swap(slot_of_pos);
	c->parent_pos_h.swap(parent_pos);
	c->level_off.swap(level_off);
	c->identity = identity;
	c->entries_dirty = true;
	invalidate_graph(c);
	return 0;
}

int hgb_upload_trs(hgb_ctx *c, uint32_t first, uint32_t count, const float *trs9) {
	if (!c)
		return fail("null context");
	if (uint64_t(first) + count > c->n_slots)
		return fail("hgb_upload_trs: range [%u,%u) outside the %u slots of the topology", first, first + count, c->n_slots);
	if (!count)
		return 0;
	CK(cudaSetDevice(c->device));
	const size_t bytes = size_t(count) * 36;
	CK(c->staging_d.reserve(bytes));
	// the caller's buffer may be pageable: the copy is then staged by the driver; pinned buffers go straight over PCIe/C2C
	CK(cudaMemcpyAsync(c->staging_d.p, trs9, bytes, cudaMemcpyHostToDevice, c->stream));
	k_scatter_trs<<<grid_for(count), 256, 0, c->stream>>>(c->staging_d.as<float>(), c->identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), first, count,
		c->trs.as<float>(), c->trs_stride);
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

int hgb_set_preset_world(hgb_ctx *c, uint32_t n, const uint32_t *slot, const float *mat12) {
	if (!c)
		return fail("null context");
	CK(cudaSetDevice(c->device));
	std::vector<uint32_t> pos(n);
	for (uint32_t i = 0; i < n; ++i) {
		if (slot[i] >= c->n_slots)
			return fail("hgb_set_preset_world: slot %u out of range", slot[i]);
		pos[i] = c->pos_of_slot[slot[i]];
	}
	if (upload(c, c->preset_pos, pos.data(), size_t(n) * 4) || upload(c, c->preset_mat, mat12, size_t(n) * 48))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	c->n_preset = n;
	c->preset_pending = n > 0;
	return 0;
}

int hgb_set_bounds(hgb_ctx *c, uint32_t n, const float *minmax6) {
	if (!c)
		return fail("null context");
	if (n >= (1u << 24))
		return fail("hgb_set_bounds: too many bounds");
	CK(cudaSetDevice(c->device));
	if (upload(c, c->bounds, minmax6, size_t(n) * 24))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	c->n_bounds = n;
	invalidate_graph(c);
	return 0;
}

int hgb_set_display_entries(hgb_ctx *c, int kind, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, const uint8_t *flags, const uint16_t *view) {
	if (!c)
		return fail("null context");
	if (kind != 0 && kind != 1)
		return fail("hgb_set_display_entries: kind must be 0 (opaque) or 1 (transparent)");
	if (n >= 0x7fffffffu)
		return fail("hgb_set_display_entries: too many entries");
	EntryTable &t = c->ent[kind];
	t.n = n;
	t.mtx_slot.assign(mtx_slot, mtx_slot + n);
	t.bounds_idx.assign(bounds_idx, bounds_idx + n);
	if (flags)
		t.flags.assign(flags, flags + n);
	else
		t.flags.assign(n, 0);
	if (view)
		t.view.assign(view, view + n);
	else
		t.view.clear();
	c->entries_dirty = true;
	invalidate_graph(c);
	return 0;
}

int hgb_set_entry_flags(hgb_ctx *c, int kind, uint32_t first, uint32_t count, const uint8_t *flags) {
	if (!c)
		return fail("null context");
	if (kind != 0 && kind != 1)
		return fail("hgb_set_entry_flags: bad kind");
	EntryTable &t = c->ent[kind];
	if (uint64_t(first) + count > t.n)
		return fail("hgb_set_entry_flags: range outside the table");
	std::copy(flags, flags + count, t.flags.begin() + first);
	c->entries_dirty = true;
	invalidate_graph(c);
	return 0;
}

int hgb_set_skins(hgb_ctx *c, uint32_t n_bones, const uint32_t *mtx_slot, const uint32_t *bind_idx, uint32_t n_bind, const float *bind_pose12) {
	if (!c)
		return fail("null context");
	CK(cudaSetDevice(c->device));
	for (uint32_t i = 0; i < n_bones; ++i)
		if (bind_idx[i] >= n_bind)
			return fail("hgb_set_skins: bind index %u out of range", bind_idx[i]);
	c->bone_mtx_slot.assign(mtx_slot, mtx_slot + n_bones);
	c->n_bones = n_bones;
	if (upload(c, c->bone_bind_idx, bind_idx, size_t(n_bones) * 4) || upload(c, c->bind_pose, bind_pose12, size_t(n_bind) * 48))
		return 1;
	CK(c->bone_out[0].reserve(std::max<size_t>(n_bones, 1) * 64));
	CK(c->bone_out[1].reserve(std::max<size_t>(n_bones, 1) * 64));
	CK(cudaStreamSynchronize(c->stream));
	c->entries_dirty = true; // bone positions are resolved with the entries
	invalidate_graph(c);
	return 0;
}

int hgb_set_anim_keys(hgb_ctx *c, uint32_t n3, const int64_t *v3_t, const float *v3_val3, const float *v3_tension, const float *v3_bias, uint32_t nq,
	const int64_t *q_t, const float *q_val4) {
	if (!c)
		return fail("null context");
	CK(cudaSetDevice(c->device));
	std::vector<float> packed(size_t(n3) * 4), bias(n3, 0.f);
	for (uint32_t i = 0; i < n3; ++i) {
		packed[size_t(i) * 4 + 0] = v3_val3[size_t(i) * 3 + 0];
		packed[size_t(i) * 4 + 1] = v3_val3[size_t(i) * 3 + 1];
		packed[size_t(i) * 4 + 2] = v3_val3[size_t(i) * 3 + 2];
		packed[size_t(i) * 4 + 3] = v3_tension ? v3_tension[i] : 0.f;
		if (v3_bias)
			bias[i] = v3_bias[i];
	}
	if (upload(c, c->v3_t, v3_t, size_t(n3) * 8) || upload(c, c->v3_val, packed.data(), size_t(n3) * 16) || upload(c, c->v3_bias, bias.data(), size_t(n3) * 4) ||
		upload(c, c->q_t, q_t, size_t(nq) * 8) || upload(c, c->q_val, q_val4, size_t(nq) * 16))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	invalidate_graph(c);
	return 0;
}

int hgb_set_tracks(hgb_ctx *c, uint32_t n, const hgb_track *tracks, uint32_t n_play) {
	if (!c)
		return fail("null context");
	for (uint32_t i = 0; i < n; ++i) {
		if (tracks[i].target_slot >= c->n_slots)
			return fail("hgb_set_tracks: track %u targets slot %u outside the topology", i, tracks[i].target_slot);
		if (tracks[i].play >= n_play || tracks[i].kind > 3)
			return fail("hgb_set_tracks: track %u has a bad play slot or kind", i);
	}
	// The reference evaluates playing anims sequentially, so when two tracks write the same TRS lane the LATER one wins
	// (scene.cpp:2383-2427). Tracks are split into rounds (k-th writer of a lane -> round k), one launch per round.
	std::vector<uint32_t> round(n, 0);
	uint32_t n_rounds = n ? 1 : 0;
	{
		std::vector<std::pair<uint64_t, uint32_t>> key(n);
		for (uint32_t i = 0; i < n; ++i) {
			const uint32_t lane = tracks[i].kind == 3 ? 1 : tracks[i].kind;
			key[i] = {(uint64_t(tracks[i].target_slot) << 2) | lane, i};
		}
		std::sort(key.begin(), key.end());
		for (uint32_t i = 1; i < n; ++i)
			if (key[i].first == key[i - 1].first) {
				round[key[i].second] = round[key[i - 1].second] + 1;
				n_rounds = std::max(n_rounds, round[key[i].second] + 1);
			}
	}
	c->tracks_h.clear();
	c->round_off.assign(1, 0);
	for (uint32_t r = 0; r < n_rounds; ++r) {
		for (uint32_t i = 0; i < n; ++i)
			if (round[i] == r)
				c->tracks_h.push_back(tracks[i]);
		c->round_off.push_back(uint32_t(c->tracks_h.size()));
	}
	c->n_tracks = n;
	c->n_play = n_play;
	CK(cudaSetDevice(c->device));
	CK(c->play_t_d.reserve(std::max<size_t>(n_play, 1) * 8));
	CK(c->play_t_h.reserve(std::max<size_t>(n_play, 1) * 8));
	c->entries_dirty = true; // target positions are resolved with the entries
	invalidate_graph(c);
	return 0;
}

int hgb_set_views(hgb_ctx *c, uint32_t n_views, const float *frustum24) {
	if (!c)
		return fail("null context");
	if (n_views > kMaxViews)
		return fail("hgb_set_views: at most %d views per pass", kMaxViews);
	CK(cudaSetDevice(c->device));
	if (c->instance_mode || c->views.n_views != int(n_views)) {
		c->entries_dirty = true; // per-view output lists are sized with the tables; n_views is baked into the graph
		invalidate_graph(c);
	}
	c->instance_mode = false;
	c->views.n_views = int(n_views);
	if (n_views) { // pageable source: staged by the runtime before returning
		float cond[kMaxViews * 24];
		precondition_planes(frustum24, n_views * 6, cond);
		CK(cudaMemcpyAsync(c->view_planes.p, cond, size_t(n_views) * 96, cudaMemcpyHostToDevice, c->stream));
	}
	return 0;
}

int hgb_set_instance_views(hgb_ctx *c, uint32_t n_inst, const float *frustum24, const uint32_t *inst_entry_start) {
	if (!c)
		return fail("null context");
	if (n_inst > 65535)
		return fail("hgb_set_instance_views: at most 65535 instances per context (uint16 view id)");
	CK(cudaSetDevice(c->device));
	const bool shape_change = !c->instance_mode || c->n_inst != n_inst;
	std::vector<float> cond(size_t(n_inst) * 24);
	precondition_planes(frustum24, n_inst * 6, cond.data());
	if (upload(c, c->inst_planes, cond.data(), size_t(n_inst) * 96))
		return 1;
	if (inst_entry_start && upload(c, c->inst_start, inst_entry_start, size_t(n_inst + 1) * 4))
		return 1;
	CK(c->inst_rank.reserve(size_t(n_inst + 1) * 4));
	CK(cudaStreamSynchronize(c->stream));
	c->instance_mode = true;
	c->n_inst = n_inst;
	c->views.n_views = 1;
	if (shape_change) {
		c->entries_dirty = true;
		invalidate_graph(c); // frusta live in device memory: new cameras alone need no re-capture
	}
	return 0;
}

} // extern "C"

// (Re)build everything that depends on both the topology and the entry tables: entry -> position maps, the per-position CSR
// used by the fused pass, bone and track target positions, compaction scratch.
static int prepare_tables(hgb_ctx *c) {
	if (!c->entries_dirty)
		return 0;
	CK(cudaSetDevice(c->device));
	const uint32_t n_pos_all = c->n_slots;
	std::vector<uint32_t> ent_off(size_t(n_pos_all) + 2, 0);
	for (int k = 0; k < 2; ++k) {
		EntryTable &t = c->ent[k];
		if (c->instance_mode && t.n && t.view.size() != t.n)
			return fail("instance views are set but entry table %d has no per-entry view ids", k);
		std::vector<uint32_t> mtx_pos(t.n), meta(t.n);
		for (uint32_t e = 0; e < t.n; ++e) {
			if (t.mtx_slot[e] >= c->n_slots)
				return fail("display entry %u (kind %d) references slot %u outside the topology", e, k, t.mtx_slot[e]);
			if (t.bounds_idx[e] >= c->n_bounds)
				return fail("display entry %u (kind %d) references bounds %u outside the table", e, k, t.bounds_idx[e]);
			mtx_pos[e] = c->pos_of_slot[t.mtx_slot[e]];
			meta[e] = t.bounds_idx[e] | (uint32_t(t.flags[e] & HGB_ENTRY_DISABLED ? 1 : 0) << 24);
			++ent_off[mtx_pos[e] + 1];
		}
		if (upload(c, t.mtx_pos, mtx_pos.data(), size_t(t.n) * 4) || upload(c, t.meta, meta.data(), size_t(t.n) * 4))
			return 1;
		if (!t.view.empty() && upload(c, t.view_d, t.view.data(), size_t(t.n) * 2))
			return 1;
		t.n_tiles = (t.n + kCompactTile - 1) / kCompactTile;
		const size_t vis_bytes = size_t(t.n_tiles + 1) * kCompactTile;
		CK(t.vis.reserve(vis_bytes));
		CK(cudaMemsetAsync(t.vis.p, 0, t.vis.cap, c->stream));
		const int n_out = c->instance_mode ? 1 : std::max(c->views.n_views, 1);
		for (int v = 0; v < n_out; ++v)
			CK(t.out[v].reserve(std::max<size_t>(t.n, 1) * 4));
		CK(t.counts.reserve(kMaxViews * 4));
		CK(cudaMemsetAsync(t.counts.p, 0, t.counts.cap, c->stream));
		CK(t.ticket.reserve(kMaxViews * 4));
		CK(t.status.reserve(size_t(std::max(t.n_tiles, 1u)) * kMaxViews * 4));
		CK(cudaStreamSynchronize(c->stream));
	}
	for (size_t i = 0; i + 1 < ent_off.size(); ++i)
		ent_off[i + 1] += ent_off[i];
	const uint32_t total = ent_off[n_pos_all];
	std::vector<uint2> pent(total);
	std::vector<uint16_t> pent_view(c->instance_mode ? total : 0);
	{
		std::vector<uint32_t> fill(ent_off.begin(), ent_off.end() - 1);
		for (int k = 0; k < 2; ++k) {
			const EntryTable &t = c->ent[k];
			for (uint32_t e = 0; e < t.n; ++e) {
				const uint32_t p = c->pos_of_slot[t.mtx_slot[e]];
				const uint32_t at = fill[p]++;
				pent[at] = make_uint2(e | (k ? kEntKindTransparent : 0u), t.bounds_idx[e] | (uint32_t(t.flags[e] & HGB_ENTRY_DISABLED ? 1 : 0) << 24));
				if (c->instance_mode)
					pent_view[at] = t.view[e];
			}
		}
	}
	if (upload(c, c->ent_off, ent_off.data(), (size_t(n_pos_all) + 1) * 4) || upload(c, c->pent, pent.data(), size_t(total) * 8))
		return 1;
	if (c->instance_mode && upload(c, c->pent_view, pent_view.data(), size_t(total) * 2))
		return 1;

	// bones and tracks: slot -> position
	std::vector<uint32_t> bone_pos(c->n_bones);
	for (uint32_t i = 0; i < c->n_bones; ++i) {
		if (c->bone_mtx_slot[i] >= c->n_slots)
			return fail("skin record %u references slot %u outside the topology", i, c->bone_mtx_slot[i]);
		bone_pos[i] = c->pos_of_slot[c->bone_mtx_slot[i]];
	}
	if (upload(c, c->bone_mtx_pos, bone_pos.data(), size_t(c->n_bones) * 4))
		return 1;
	std::vector<Track> tr(c->tracks_h.size());
	for (size_t i = 0; i < tr.size(); ++i) {
		const hgb_track &h = c->tracks_h[i];
		if (h.target_slot >= c->n_slots)
			return fail("track %zu targets slot %u outside the topology", i, h.target_slot);
		tr[i] = Track{c->pos_of_slot[h.target_slot], h.key_off, h.key_count, h.play, h.kind, 0};
	}
	if (upload(c, c->tracks_d, tr.data(), tr.size() * sizeof(Track)))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	c->entries_dirty = false;
	invalidate_graph(c);
	return 0;
}
