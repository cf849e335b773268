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

// hg::MinMax {mn.xyz, mx.xyz} (foundation/minmax.h:13-18) -> device form {mn.x, mx.x, mn.y, mx.y, mn.z, mx.z} (see minmax_transform)
static void interleave_bounds(const float *in, size_t n, std::vector<float> &out) {
	out.resize(std::max<size_t>(n, 1) * 6);
	for (size_t i = 0; i < n; ++i)
		for (int k = 0; k < 3; ++k) {
			out[i * 6 + 2 * k] = in[i * 6 + k];
			out[i * 6 + 2 * k + 1] = in[i * 6 + 3 + k];
		}
}

// conditioned planes -> the expanded form of survives_frustum_x: {x, |x|, y, |y|} {z, |z|, -2w, 0}
static void expand_planes(const float *cond, uint32_t n_planes, float *out) {
	for (uint32_t i = 0; i < n_planes; ++i) {
		const float *p = cond + i * 4;
		float *o = out + i * 8;
		o[0] = p[0], o[1] = fabsf(p[0]), o[2] = p[1], o[3] = fabsf(p[1]);
		o[4] = p[2], o[5] = fabsf(p[2]), o[6] = p[3], o[7] = 0.f;
	}
}

static inline dim3 grid_for(uint32_t n, uint32_t block = 256) { return dim3((n + block - 1) / block); }

static int prepare_tables(hgb_ctx *c);

// "Enable" tracks are evaluated on the device (k_sample_enable): before the host-side tables are rebuilt or edited, the state of the
// tracked nodes is read back and folded into them, so that a rebuild does not resurrect a stale flag.
static int fold_enable_state(hgb_ctx *c) {
	if (!c->en_state_dirty)
		return 0;
	c->en_state_dirty = false;
	const uint32_t n = uint32_t(c->en_target_slot_h.size());
	if (!n || !c->en_resolved)
		return 0;
	CK(cudaSetDevice(c->device));
	CK(c->en_scratch.reserve(n));
	k_gather_disabled<<<grid_for(n, 128), 128, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->en_target_pos.as<uint32_t>(), n, c->en_scratch.as<uint8_t>());
	++c->launches;
	std::vector<uint8_t> st(n);
	CK(cudaMemcpyAsync(st.data(), c->en_scratch.p, n, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	c->en_disabled_h = st;
	c->en_known_h.assign(n, 1);
	for (int k = 0; k < 2; ++k) { // entries of the tracked nodes take the node's state
		EntryTable &t = c->ent[k];
		for (uint32_t e = 0; e < t.n; ++e) {
			const auto it = std::lower_bound(c->en_target_slot_h.begin(), c->en_target_slot_h.end(), t.mtx_slot[e]);
			if (it != c->en_target_slot_h.end() && *it == t.mtx_slot[e]) {
				const uint8_t dis = st[it - c->en_target_slot_h.begin()];
				t.flags[e] = uint8_t(dis ? (t.flags[e] | HGB_ENTRY_DISABLED) : (t.flags[e] & ~HGB_ENTRY_DISABLED));
			}
		}
	}
	return 0;
}

// the hierarchy as it stands (parents by slot, used flags) — the input of a full depth sort after an edit the fast paths cannot take
static void snapshot_parents(const hgb_ctx *c, std::vector<uint32_t> &parent, std::vector<uint8_t> &used) {
	parent.assign(c->n_slots, HGB_ROOT);
	used.assign(c->n_slots, 0);
	const uint32_t n_all = c->append_slack ? c->cap_slots : c->n_pos;
	for (uint32_t p = 0; p < n_all; ++p) {
		if (c->append_slack && !c->live_h[p])
			continue;
		const uint32_t s_ = c->slot_of_pos[p];
		if (s_ >= c->n_slots)
			continue;
		used[s_] = 1;
		parent[s_] = c->parent_pos_h[p] == kRoot ? HGB_ROOT : c->slot_of_pos[c->parent_pos_h[p]];
	}
}
static inline bool pos_is_live(const hgb_ctx *c, uint32_t pos) { return c->append_slack ? (pos < c->cap_slots && c->live_h[pos]) : pos < c->n_pos; }
// slots a range call may name: the topology's, or — with append slack — every slot the device arrays hold (rows of free slots are scratch)
static inline uint32_t slot_limit(const hgb_ctx *c) { return std::max(c->n_slots, c->cap_slots); }

extern "C" {

int hgb_set_topology(hgb_ctx *c, uint32_t n, const uint32_t *parent_slot, const uint8_t *used) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (fold_enable_state(c)) // the tiles are about to be rebuilt
		return 1;
	++c->rebuilds;
	c->en_resolved = false;
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
	// in-place edits (hgb_options.append_slack): the holes and `slack` further positions behind them form the append region — free
	// positions (parent row kDead) that hgb_append_nodes hands out; slots >= n map to positions of the same index
	const uint32_t slack = c->append_slack, cap = n + slack;
	if (cap >= 0x00ffffffu)
		return fail("hgb_set_topology: %u slots + %u positions of append slack exceed 2^24-1", n, slack);
	for (uint32_t s = n; s < cap; ++s)
		slot_of_pos.push_back(s);

	std::vector<uint32_t> pos_of_slot(cap);
	bool identity = true;
	for (uint32_t p = 0; p < cap; ++p) {
		pos_of_slot[slot_of_pos[p]] = p;
		identity = identity && slot_of_pos[p] == p;
	}
	std::vector<uint32_t> parent_pos(cap, slack ? kDead : kRoot);
	for (uint32_t p = 0; p < n_pos; ++p) {
		const uint32_t par = parent_of(slot_of_pos[p]);
		parent_pos[p] = par == HGB_ROOT ? kRoot : pos_of_slot[par];
	}

	// carry the device TRS over to the new order (animation may have modified it on the device)
	const uint32_t keep = std::min(c->n_slots, n);
	const bool carry = keep > 0 && c->tiles.p;
	if (carry) {
		CK(c->staging_d.reserve(size_t(keep) * 9 * 4));
		k_gather_trs<<<grid_for(keep), 256, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), 0, keep,
			c->staging_d.as<float>(), c->anim_val.as<float4>());
		++c->launches;
		CK(cudaStreamSynchronize(c->stream));
	}
	// ... and both world buffers: transform_worlds / previous_transform_worlds are indexed by slot in the reference and survive hierarchy
	// edits (scene.cpp:280-304), so the frame after an edit still sees the right previous-frame matrices (motion vectors, bone twins).
	// Nothing to do when old and new order are both the slot order and the buffers do not grow.
	DBuf carry_w[2];
	const bool grow = c->world[0].cap < std::max<size_t>(cap, 1) * 48 || c->world[1].cap < std::max<size_t>(cap, 1) * 48;
	const bool carry_worlds = carry && c->world[0].p && c->world[1].p && !(c->identity && identity && !grow);
	if (carry_worlds) {
		for (int i = 0; i < 2; ++i) {
			CK(carry_w[i].reserve(size_t(keep) * 48));
			k_gather_worlds<<<grid_for(keep * 3), 256, 0, c->stream>>>(c->world[i].as<float4>(), c->identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), 0, keep,
				carry_w[i].as<float4>());
			++c->launches;
		}
		CK(cudaStreamSynchronize(c->stream));
	}

	const uint32_t n_tiles = (cap + 31) / 32;
	DBuf new_tiles, rows;
	CK(new_tiles.reserve(std::max<size_t>(n_tiles, 1) * kTileWords * 4));
	CK(cudaMemsetAsync(new_tiles.p, 0, new_tiles.cap, c->stream));
	if (upload(c, c->pos_of_slot_d, pos_of_slot.data(), size_t(cap) * 4) || upload(c, rows, parent_pos.data(), size_t(cap) * 4))
		return 1;
	if (cap) {
		k_fill_rows<<<grid_for(cap), 256, 0, c->stream>>>(new_tiles.as<uint32_t>(), kRowParent, rows.as<uint32_t>(), cap);
		++c->launches;
	}
	if (carry) {
		k_scatter_trs<<<grid_for(keep), 256, 0, c->stream>>>(c->staging_d.as<float>(), identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), 0, keep,
			new_tiles.as<uint32_t>(), nullptr); // the new tiles hold plain values: triples are re-bound by prepare_tables
		++c->launches;
	}
	CK(cudaStreamSynchronize(c->stream)); // host vectors above go out of scope
	rows.release();
	c->tiles.release();
	c->tiles = new_tiles;
	c->n_tiles = n_tiles;
	c->n_bound = 0;
	for (int i = 0; i < 2; ++i) {
		const size_t bytes = std::max<size_t>(cap, 1) * 48;
		if (c->world[i].cap < bytes) {
			CK(c->world[i].reserve(bytes));
			CK(cudaMemsetAsync(c->world[i].p, 0, c->world[i].cap, c->stream)); // new slots: value-initialised rows, like std::vector::resize
		}
		if (carry_worlds) {
			k_scatter_worlds<<<grid_for(keep * 3), 256, 0, c->stream>>>(carry_w[i].as<float4>(), identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), keep,
				c->world[i].as<float4>());
			++c->launches;
		}
	}
	if (carry_worlds) {
		CK(cudaStreamSynchronize(c->stream));
		carry_w[0].release();
		carry_w[1].release();
	}
	CK(c->preset.reserve(std::max<size_t>(cap, 4)));
	CK(cudaMemsetAsync(c->preset.p, 0, c->preset.cap, c->stream));
	c->preset_dirty = false;

	c->cap_slots = cap;
	if (slack) { // host mirror of what the in-place edits need to know
		c->live_h.assign(cap, 0);
		std::fill(c->live_h.begin(), c->live_h.begin() + n_pos, uint8_t(1));
		c->late_depth_h.assign(cap, 0);
		c->child_cnt_h.assign(cap, 0);
		for (uint32_t s = 0; s < n; ++s)
			if (is_used(s))
				c->child_cnt_h[pos_of_slot[s]] = child_cnt[s + 1] - child_cnt[s];
		c->has_entry_h.assign(cap, 0);
	}
	c->n_slots = n;
	c->n_pos = n_pos;
	c->pos_of_slot.swap(pos_of_slot);
	c->slot_of_pos.swap(slot_of_pos);
	c->parent_pos_h.swap(parent_pos);
	c->level_off.swap(level_off);
	c->identity = identity;
	c->entries_dirty = true;
	c->obj_dirty = true;
	invalidate_graph(c);
	return 0;
}

// Incremental hierarchy edit: new parents for a few transforms. When every moved transform stays in its hierarchy LEVEL (its new parent
// sits one level above it, or it was and stays a root) the depth-sorted layout is still valid: only the parent row of those positions
// is rewritten — a few bytes, no re-sort, no table rebuild, no graph re-capture (parent rows are data the kernels read). A same-level
// move cannot close a cycle (the new parent is not below the node). Anything else falls back to the full hgb_set_topology rebuild.
int hgb_set_parents(hgb_ctx *c, uint32_t n, const uint32_t *slot, const uint32_t *parent_slot, int *out_rebuilt) {
	if (!c)
		return fail("null context");
	if (out_rebuilt)
		*out_rebuilt = 0;
	if (!n)
		return 0;
	join_post(c);
	CK(cudaSetDevice(c->device));
	auto level_of = [&](uint32_t pos) { return uint32_t(std::upper_bound(c->level_off.begin(), c->level_off.end(), pos) - c->level_off.begin()) - 1u; };
	struct Undo {
		uint32_t pos, par;
		uint8_t depth;
	};
	std::vector<Undo> undo; // host mirror of the append region: edits are applied one by one and rolled back when the batch needs a rebuild
	std::vector<uint32_t> pos(n), par(n);
	bool fast = true;
	for (uint32_t i = 0; i < n; ++i) {
		if (slot[i] >= c->n_slots || !pos_is_live(c, c->pos_of_slot[slot[i]]))
			return fail("hgb_set_parents: slot %u is not a live transform", slot[i]);
		pos[i] = c->pos_of_slot[slot[i]];
		const uint32_t ps = parent_slot[i];
		const bool root = ps == HGB_ROOT || ps >= c->n_slots || !pos_is_live(c, c->pos_of_slot[ps]) || ps == slot[i]; // stale / unused parent = root (scene.cpp:268-272)
		par[i] = root ? kRoot : c->pos_of_slot[ps];
		if (!fast)
			continue;
		uint8_t depth = 0;
		if (pos[i] >= c->n_pos) { // a node of the append region: anywhere, as long as it has no children and its chain stays short
			if (c->child_cnt_h[pos[i]])
				fast = false;
			else if (!root && par[i] >= c->n_pos) {
				if (c->late_depth_h[par[i]] + 1 >= kMaxWalk)
					fast = false;
				depth = uint8_t(c->late_depth_h[par[i]] + 1);
			}
		} else { // sorted region: the node must stay in its level
			const uint32_t lv = level_of(pos[i]);
			if (root ? lv != 0 : (par[i] >= c->n_pos || level_of(par[i]) + 1 != lv))
				fast = false;
		}
		if (fast && c->append_slack) {
			undo.push_back({pos[i], c->parent_pos_h[pos[i]], c->late_depth_h[pos[i]]});
			if (c->parent_pos_h[pos[i]] != kRoot)
				--c->child_cnt_h[c->parent_pos_h[pos[i]]];
			if (!root)
				++c->child_cnt_h[par[i]];
			c->parent_pos_h[pos[i]] = par[i];
			c->late_depth_h[pos[i]] = depth;
		}
	}
	if (!fast) { // levels change: full rebuild from the current parents with the edits applied
		for (size_t k = undo.size(); k-- > 0;) {
			const uint32_t now = c->parent_pos_h[undo[k].pos];
			if (now != kRoot)
				--c->child_cnt_h[now];
			if (undo[k].par != kRoot)
				++c->child_cnt_h[undo[k].par];
			c->parent_pos_h[undo[k].pos] = undo[k].par;
			c->late_depth_h[undo[k].pos] = undo[k].depth;
		}
		std::vector<uint32_t> parent;
		std::vector<uint8_t> used;
		snapshot_parents(c, parent, used);
		for (uint32_t i = 0; i < n; ++i)
			parent[slot[i]] = parent_slot[i];
		if (out_rebuilt)
			*out_rebuilt = 1;
		return hgb_set_topology(c, c->n_slots, parent.data(), used.data());
	}
	if (!c->append_slack)
		for (uint32_t i = 0; i < n; ++i)
			c->parent_pos_h[pos[i]] = par[i];
	CK(c->staging.reserve(size_t(n) * 8));
	uint32_t *st = c->staging.as<uint32_t>();
	CK(cudaStreamSynchronize(c->stream)); // the pinned staging block may still feed an earlier copy
	memcpy(st, pos.data(), size_t(n) * 4);
	memcpy(st + n, par.data(), size_t(n) * 4);
	CK(c->staging_d.reserve(size_t(n) * 8));
	CK(cudaMemcpyAsync(c->staging_d.p, st, size_t(n) * 8, cudaMemcpyHostToDevice, c->stream));
	k_set_parent_rows<<<grid_for(n, 128), 128, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->staging_d.as<uint32_t>(), c->staging_d.as<uint32_t>() + n, n);
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

// small host table -> device scratch on the context's stream (the copy is staged by the driver before the call returns: the caller's
// vector may go out of scope)
static int edit_upload(hgb_ctx *c, DBuf &dst, const void *src, size_t bytes) {
	CK(dst.reserve(bytes));
	CK(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
	return 0;
}

// New LEAF transforms (Scene::CreateNode + CreateTransform + SetParent of a fresh node, scene.cpp:92-137, component.cpp:184-193) without
// a depth sort: each takes a free position of the append region, which k_update_top computes behind the last level by walking up to
// the first ancestor in the sorted part. Falls back to the full rebuild when the context has no append slack, the region is full, a
// slot is out of range or a chain inside the region would exceed kMaxWalk.
int hgb_append_nodes(hgb_ctx *c, uint32_t n, const uint32_t *slot, const uint32_t *parent_slot, const float *trs9, int *out_rebuilt) {
	if (!c)
		return fail("null context");
	if (out_rebuilt)
		*out_rebuilt = 0;
	if (!n)
		return 0;
	join_post(c);
	CK(cudaSetDevice(c->device));
	std::vector<NodeWrite> w;
	w.reserve(n);
	auto level_of = [&](uint32_t pos) { return uint32_t(std::upper_bound(c->level_off.begin(), c->level_off.end(), pos) - c->level_off.begin()) - 1u; };
	bool fast = c->append_slack && c->tiles.p;
	uint32_t n_slots = c->n_slots;
	for (uint32_t i = 0; i < n && fast; ++i) {
		if (slot[i] >= c->cap_slots) {
			fast = false;
			break;
		}
		const uint32_t p = c->pos_of_slot[slot[i]];
		if (c->live_h[p]) {
			for (const NodeWrite &x : w) // roll the batch back before reporting
				c->live_h[x.pos] = 0;
			return fail("hgb_append_nodes: slot %u already holds a live transform", slot[i]);
		}
		const uint32_t ps = parent_slot ? parent_slot[i] : HGB_ROOT;
		const bool root = ps == HGB_ROOT || ps >= c->cap_slots || !c->live_h[c->pos_of_slot[ps]] || ps == slot[i];
		const uint32_t par = root ? kRoot : c->pos_of_slot[ps];
		uint8_t depth = 0;
		if (p < c->n_pos) {
			// the slot of a node removed from the SORTED part: its position belongs to a level, so the new node may only hang below a
			// node of an earlier level (whose row is final when this level runs) or be a root
			if (!root && (par >= c->n_pos || level_of(par) >= level_of(p))) {
				fast = false;
				break;
			}
		} else if (!root && par >= c->n_pos) {
			if (c->late_depth_h[par] + 1 >= kMaxWalk) {
				fast = false;
				break;
			}
			depth = uint8_t(c->late_depth_h[par] + 1);
		}
		NodeWrite x{};
		x.pos = p, x.parent = par, x.zero_world = 1;
		memcpy(x.trs, trs9 + size_t(i) * 9, 36);
		w.push_back(x);
		c->live_h[p] = 1; // a later node of the batch may hang below this one
		c->late_depth_h[p] = depth;
		n_slots = std::max(n_slots, slot[i] + 1);
	}
	if (!fast) {
		for (const NodeWrite &x : w)
			c->live_h[x.pos] = 0;
		std::vector<uint32_t> parent;
		std::vector<uint8_t> used;
		snapshot_parents(c, parent, used);
		uint32_t n_new = c->n_slots;
		for (uint32_t i = 0; i < n; ++i)
			n_new = std::max(n_new, slot[i] + 1);
		parent.resize(n_new, HGB_ROOT);
		used.resize(n_new, 0);
		for (uint32_t i = 0; i < n; ++i) {
			if (used[slot[i]])
				return fail("hgb_append_nodes: slot %u already holds a live transform", slot[i]);
			used[slot[i]] = 1;
			parent[slot[i]] = parent_slot ? parent_slot[i] : HGB_ROOT;
		}
		if (out_rebuilt)
			*out_rebuilt = 1;
		if (hgb_set_topology(c, n_new, parent.data(), used.data()))
			return 1;
		for (uint32_t i = 0; i < n; ++i)
			if (hgb_upload_trs(c, slot[i], 1, trs9 + size_t(i) * 9))
				return 1;
		return 0;
	}
	for (const NodeWrite &x : w) {
		c->parent_pos_h[x.pos] = x.parent;
		c->child_cnt_h[x.pos] = 0;
		c->has_entry_h[x.pos] = 0;
		if (x.parent != kRoot)
			++c->child_cnt_h[x.parent];
	}
	c->n_slots = n_slots;
	if (edit_upload(c, c->edit_d, w.data(), w.size() * sizeof(NodeWrite)))
		return 1;
	k_write_nodes<<<grid_for(uint32_t(w.size()), 128), 128, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->world[0].as<float>(), c->world[1].as<float>(),
		c->preset.as<uint8_t>(), c->edit_d.as<NodeWrite>(), uint32_t(w.size()));
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

// Remove transforms (Scene::DestroyNode / DestroyTransform + GarbageCollect, scene.cpp:139-237). A childless node goes in place: its
// display entries are disabled on the device, a position of the append region becomes free again, a position of the sorted part idles as a
// childless root until the next depth sort. A node that still has children re-roots them (scene.cpp:268-272: a stale parent ref is no
// parent), which changes their levels: full rebuild.
int hgb_remove_nodes(hgb_ctx *c, uint32_t n, const uint32_t *slot, int *out_rebuilt) {
	if (!c)
		return fail("null context");
	if (out_rebuilt)
		*out_rebuilt = 0;
	if (!n)
		return 0;
	join_post(c);
	CK(cudaSetDevice(c->device));
	for (uint32_t i = 0; i < n; ++i)
		if (slot[i] >= c->n_slots || !pos_is_live(c, c->pos_of_slot[slot[i]]))
			return fail("hgb_remove_nodes: slot %u is not a live transform", slot[i]);
	for (uint32_t i = 0; i < n; ++i) // the display entries these nodes own today are dead from here on (prepare_tables)
		c->tomb[slot[i]] = {c->ent[0].n, c->ent[1].n};
	bool fast = c->append_slack != 0;
	std::vector<uint32_t> pos;
	if (fast) { // children removed earlier in the batch no longer count
		std::vector<uint32_t> cnt_undo;
		for (uint32_t i = 0; i < n && fast; ++i) {
			const uint32_t p = c->pos_of_slot[slot[i]];
			if (c->child_cnt_h[p] || !c->live_h[p]) {
				fast = false;
				break;
			}
			pos.push_back(p);
			c->live_h[p] = 0;
			if (c->parent_pos_h[p] != kRoot)
				--c->child_cnt_h[c->parent_pos_h[p]];
		}
		if (!fast)
			for (uint32_t p : pos) {
				c->live_h[p] = 1;
				if (c->parent_pos_h[p] != kRoot)
					++c->child_cnt_h[c->parent_pos_h[p]];
			}
	}
	if (!fast) {
		std::vector<uint32_t> parent;
		std::vector<uint8_t> used;
		snapshot_parents(c, parent, used);
		for (uint32_t i = 0; i < n; ++i)
			used[slot[i]] = 0;
		if (out_rebuilt)
			*out_rebuilt = 1;
		return hgb_set_topology(c, c->n_slots, parent.data(), used.data());
	}
	for (uint32_t p : pos) {
		c->parent_pos_h[p] = p >= c->n_pos ? kDead : kRoot;
		c->late_depth_h[p] = 0;
		c->has_entry_h[p] = 0; // (a retired position of the sorted part keeps a disabled entry in its row until a new node overwrites it)
	}
	if (edit_upload(c, c->edit_d, pos.data(), pos.size() * 4))
		return 1;
	RetireDev d{};
	d.tiles = c->tiles.as<uint32_t>();
	if (!c->entries_dirty) // else the tables are rebuilt from the host arrays before the next frame (entries of dead slots come out disabled)
		for (int k = 0; k < 2; ++k) {
			d.ent_meta[k] = c->ent[k].meta.as<uint32_t>();
			d.vis[k][0] = c->ent[k].vis[0].as<uint8_t>();
			d.vis[k][1] = c->ent[k].vis[1].as<uint8_t>();
		}
	d.ent_off = c->ent_off.as<uint32_t>();
	d.pent = c->pent.as<uint2>();
	d.pos = c->edit_d.as<uint32_t>();
	d.n = uint32_t(pos.size());
	d.dead_from = c->n_pos;
	k_retire_nodes<<<grid_for(d.n, 128), 128, 0, c->stream>>>(d);
	++c->launches;
	CK(cudaGetLastError());
	// the host arrays follow, so that a later table rebuild does not resurrect the entries (prepare_tables disables entries of dead slots)
	return 0;
}

int hgb_upload_trs(hgb_ctx *c, uint32_t first, uint32_t count, const float *trs9) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (uint64_t(first) + count > slot_limit(c))
		return fail("hgb_upload_trs: range [%u,%u) outside the %u slots of the topology", first, first + count, slot_limit(c));
	if (!count)
		return 0;
	CK(cudaSetDevice(c->device));
	const size_t bytes = size_t(count) * 36;
	CK(c->staging_d.reserve(bytes));
	// the caller's buffer may be pageable: the copy is then staged by the driver; pinned buffers go straight over PCIe/C2C
	CK(cudaMemcpyAsync(c->staging_d.p, trs9, bytes, cudaMemcpyHostToDevice, c->stream));
	k_scatter_trs<<<grid_for(count), 256, 0, c->stream>>>(c->staging_d.as<float>(), c->identity ? nullptr : c->pos_of_slot_d.as<uint32_t>(), first, count,
		c->tiles.as<uint32_t>(), c->anim_val.as<float4>());
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

int hgb_set_preset_world(hgb_ctx *c, uint32_t n, const uint32_t *slot, const float *mat12) {
	if (!c)
		return fail("null context");
	join_post(c);
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

// Scene::SetTransformWorldMatrix (is_local = 0, component.cpp:265-281) / SetTransformLocalMatrix (is_local = 1, :236-248) for a batch of
// transforms, with the reference's sequential semantics: the batch is cut wherever an entry touches a row an earlier entry of the same
// launch reads or writes.
int hgb_set_matrices(hgb_ctx *c, uint32_t n, const uint32_t *slot, const float *mat12, int is_local) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (!n)
		return 0;
	CK(cudaSetDevice(c->device));
	if (prepare_tables(c)) // triples bound to tracks must be known before TRS rows are written
		return 1;
	std::vector<uint32_t> pos(n);
	for (uint32_t i = 0; i < n; ++i) {
		if (slot[i] >= c->n_slots || !pos_is_live(c, c->pos_of_slot[slot[i]]))
			return fail("hgb_set_matrices: slot %u is not a live transform", slot[i]);
		pos[i] = c->pos_of_slot[slot[i]];
	}
	if (upload(c, c->preset_pos, pos.data(), size_t(n) * 4) || upload(c, c->preset_mat, mat12, size_t(n) * 48))
		return 1;
	c->preset_pending = false; // the staging buffers are shared with hgb_set_preset_world
	c->n_preset = 0;
	std::vector<uint32_t> &stamp_w = c->stamp_w, &stamp_r = c->stamp_r; // run id in which a position was written / read as a parent
	stamp_w.assign(c->n_slots, 0);
	stamp_r.assign(c->n_slots, 0);
	uint32_t run = 1, first = 0;
	auto flush = [&](uint32_t end) -> int {
		if (end > first) {
			k_set_matrices<<<grid_for(end - first, 128), 128, 0, c->stream>>>(c->preset_pos.as<uint32_t>() + first, c->preset_mat.as<float4>() + size_t(first) * 3,
				end - first, is_local, c->world[c->cur].as<float>(), c->tiles.as<uint32_t>(), c->anim_val.as<float4>());
			++c->launches;
			CK(cudaGetLastError());
		}
		first = end;
		++run;
		return 0;
	};
	for (uint32_t i = 0; i < n; ++i) {
		const uint32_t p = pos[i], par = c->parent_pos_h[p];
		const bool clash = stamp_w[p] == run || stamp_r[p] == run || (par != kRoot && stamp_w[par] == run);
		if (clash && flush(i))
			return 1;
		stamp_w[p] = run;
		if (par != kRoot)
			stamp_r[par] = run;
	}
	if (flush(n))
		return 1;
	CK(cudaStreamSynchronize(c->stream)); // pos[] and the caller's matrices are free again
	return 0;
}

int hgb_set_objects(hgb_ctx *c, uint32_t n_obj, const uint32_t *mtx_slot, const uint32_t *obounds_idx, uint32_t n_obounds, const float *obounds6,
	const uint32_t *seg, uint32_t n_seg) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (n_obounds >= (1u << 24))
		return fail("hgb_set_objects: too many object bounds");
	if (n_obj && n_seg == 0)
		return fail("hgb_set_objects: at least one segment is needed");
	for (uint32_t i = 0; i < n_obj; ++i) {
		if (obounds_idx[i] >= n_obounds)
			return fail("hgb_set_objects: object %u references bounds %u outside the table", i, obounds_idx[i]);
		if (seg && seg[i] >= n_seg)
			return fail("hgb_set_objects: object %u is in segment %u of %u", i, seg[i], n_seg);
	}
	CK(cudaSetDevice(c->device));
	c->obj_slot_h.assign(mtx_slot, mtx_slot + n_obj);
	c->obj_ob_h.assign(obounds_idx, obounds_idx + n_obj);
	if (seg)
		c->obj_seg_h.assign(seg, seg + n_obj);
	else
		c->obj_seg_h.clear();
	c->n_obj = n_obj, c->n_obounds = n_obounds, c->n_seg = n_seg;
	std::vector<float> il;
	interleave_bounds(obounds6, n_obounds, il);
	if (upload(c, c->obounds, il.data(), size_t(n_obounds) * 24))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	c->obj_dirty = true;
	return 0;
}

int hgb_set_bounds(hgb_ctx *c, uint32_t n, const float *minmax6) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (n >= (1u << 24))
		return fail("hgb_set_bounds: too many bounds");
	CK(cudaSetDevice(c->device));
	interleave_bounds(minmax6, n, c->bounds_stage);
	if (upload(c, c->bounds, c->bounds_stage.data(), size_t(n) * 24))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	c->n_bounds = n;
	invalidate_graph(c);
	return 0;
}

int hgb_set_display_entries(hgb_ctx *c, int kind, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, const uint8_t *flags, const uint16_t *view) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (kind != 0 && kind != 1)
		return fail("hgb_set_display_entries: kind must be 0 (opaque) or 1 (transparent)");
	if (n >= 0x7fffffffu)
		return fail("hgb_set_display_entries: too many entries");
	EntryTable &t = c->ent[kind];
	t.n = n;
	t.n_dev = n + c->append_slack; // free entries behind the table for hgb_append_display_entries
	t.n_records = 0;
	for (auto it = c->tomb.begin(); it != c->tomb.end();) { // a new table: nothing in it belongs to a removed node
		it->second[kind] = 0;
		it = (it->second[0] | it->second[1]) ? std::next(it) : c->tomb.erase(it);
	}
	if (c->append_slack) { // appended entries must not re-allocate 16M-element host arrays
		t.mtx_slot.reserve(t.n_dev), t.bounds_idx.reserve(t.n_dev), t.flags.reserve(t.n_dev);
		if (view)
			t.view.reserve(t.n_dev);
	}
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
	join_post(c);
	if (fold_enable_state(c)) // this explicit write must land AFTER what the tracks did so far
		return 1;

	if (kind != 0 && kind != 1)
		return fail("hgb_set_entry_flags: bad kind");
	EntryTable &t = c->ent[kind];
	if (uint64_t(first) + count > t.n)
		return fail("hgb_set_entry_flags: range outside the table");
	std::copy(flags, flags + count, t.flags.begin() + first);
	if (!count || c->entries_dirty) // the tables are rebuilt from the host arrays before the next frame anyway
		return 0;
	// tables are live: flip the disabled bit where the kernels read it (entry table, the node's tile row or its overflow record) —
	// EnableNode / DisableNode of a few nodes costs a few bytes, no table rebuild and no graph capture
	CK(cudaSetDevice(c->device));
	std::vector<uint8_t> dis(count);
	for (uint32_t i = 0; i < count; ++i)
		dis[i] = (flags[i] & HGB_ENTRY_DISABLED) ? 1 : 0;
	if (edit_upload(c, c->edit_d, dis.data(), count))
		return 1;
	k_set_entry_flags<<<grid_for(count, 128), 128, 0, c->stream>>>(c->tiles.as<uint32_t>(), t.mtx_pos.as<uint32_t>(), t.meta.as<uint32_t>(), c->ent_off.as<uint32_t>(),
		c->pent.as<uint2>(), c->edit_d.as<uint8_t>(), first, count, kind ? kEntKindTransparent : 0u);
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

// New display entries behind the last one of a table (a node created at the end of the scene's node list lands at the end of
// GetModelDisplayLists' output, scene.cpp:318-404). In place when the context has append slack, the tables are live, the views are
// shared (no per-instance segments) and each entry is the first one of its node; else the tables are rebuilt before the next frame.
int hgb_append_display_entries(hgb_ctx *c, int kind, uint32_t n, const uint32_t *mtx_slot, const uint32_t *bounds_idx, const uint8_t *flags, const uint16_t *view,
	int *out_rebuilt) {
	if (!c)
		return fail("null context");
	if (out_rebuilt)
		*out_rebuilt = 0;
	if (!n)
		return 0;
	join_post(c);
	if (kind != 0 && kind != 1)
		return fail("hgb_append_display_entries: kind must be 0 (opaque) or 1 (transparent)");
	EntryTable &t = c->ent[kind];
	if (uint64_t(t.n) + n >= 0x7fffffffu)
		return fail("hgb_append_display_entries: too many entries");
	if ((view != nullptr) != !t.view.empty() && t.n)
		return fail("hgb_append_display_entries: per-entry view ids must be given for all entries of a table or for none");
	for (uint32_t i = 0; i < n; ++i) {
		if (mtx_slot[i] >= c->n_slots)
			return fail("hgb_append_display_entries: entry %u references slot %u outside the topology", i, mtx_slot[i]);
		if (bounds_idx[i] >= c->n_bounds)
			return fail("hgb_append_display_entries: entry %u references bounds %u outside the table", i, bounds_idx[i]);
	}
	const uint32_t first = t.n;
	bool fast = c->append_slack && !c->entries_dirty && !c->instance_mode && !view && uint64_t(first) + n <= t.n_dev && t.meta.p;
	std::vector<EntryWrite> w(fast ? n : 0);
	std::vector<uint32_t> marked;
	for (uint32_t i = 0; i < n && fast; ++i) {
		const uint32_t p = c->pos_of_slot[mtx_slot[i]];
		if (!c->live_h[p] || c->has_entry_h[p]) { // a second entry of a node lives in the overflow CSR: table rebuild
			fast = false;
			break;
		}
		c->has_entry_h[p] = 1;
		marked.push_back(p);
		w[i] = EntryWrite{first + i, p, bounds_idx[i] | ((flags && (flags[i] & HGB_ENTRY_DISABLED)) ? kEntDisabled : 0u)};
	}
	t.mtx_slot.insert(t.mtx_slot.end(), mtx_slot, mtx_slot + n);
	t.bounds_idx.insert(t.bounds_idx.end(), bounds_idx, bounds_idx + n);
	if (flags)
		t.flags.insert(t.flags.end(), flags, flags + n);
	else
		t.flags.insert(t.flags.end(), n, uint8_t(0));
	if (view)
		t.view.insert(t.view.end(), view, view + n);
	t.n = first + n;
	if (!fast) {
		for (uint32_t p : marked)
			c->has_entry_h[p] = 0;
		t.n_dev = std::max(t.n_dev, t.n + c->append_slack);
		c->entries_dirty = true;
		invalidate_graph(c);
		if (out_rebuilt)
			*out_rebuilt = 1;
		return 0;
	}
	CK(cudaSetDevice(c->device));
	if (edit_upload(c, c->edit_d, w.data(), w.size() * sizeof(EntryWrite)))
		return 1;
	k_append_entries<<<grid_for(n, 128), 128, 0, c->stream>>>(c->tiles.as<uint32_t>(), t.mtx_pos.as<uint32_t>(), t.meta.as<uint32_t>(), t.vis[0].as<uint8_t>(),
		t.vis[1].as<uint8_t>(), c->edit_d.as<EntryWrite>(), n, kind ? kEntKindTransparent : 0u);
	++c->launches;
	CK(cudaGetLastError());
	return 0;
}

int hgb_set_skins(hgb_ctx *c, uint32_t n_bones, const uint32_t *mtx_slot, const uint32_t *bind_idx, uint32_t n_bind, const float *bind_pose12) {
	if (!c)
		return fail("null context");
	join_post(c);
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
	join_post(c);
	CK(cudaSetDevice(c->device));
	// one 32-byte record per key: {t (8), v.x, v.y | v.z, tension or v.w, bias, 0}; quaternion keys after the vec3 keys
	const size_t total = size_t(n3) + nq;
	CK(c->staging.reserve((total + 4) * 32));
	uint32_t *rec = c->staging.as<uint32_t>();
	auto f2u = [](float f) { uint32_t u; memcpy(&u, &f, 4); return u; };
	for (size_t i = 0; i < n3; ++i) {
		uint32_t *r = rec + i * 8;
		memcpy(r, v3_t + i, 8);
		r[2] = f2u(v3_val3[i * 3 + 0]), r[3] = f2u(v3_val3[i * 3 + 1]), r[4] = f2u(v3_val3[i * 3 + 2]);
		r[5] = f2u(v3_tension ? v3_tension[i] : 0.f), r[6] = f2u(v3_bias ? v3_bias[i] : 0.f), r[7] = 0;
	}
	for (size_t i = 0; i < nq; ++i) {
		uint32_t *r = rec + (size_t(n3) + i) * 8;
		memcpy(r, q_t + i, 8);
		r[2] = f2u(q_val4[i * 4 + 0]), r[3] = f2u(q_val4[i * 4 + 1]), r[4] = f2u(q_val4[i * 4 + 2]), r[5] = f2u(q_val4[i * 4 + 3]);
		r[6] = r[7] = 0;
	}
	// 4 keys of zero padding after the pool: k_sample_tracks reads fixed 4-key windows (128 contiguous bytes) starting at any key
	memset(rec + total * 8, 0, 4 * 32);
	if (upload(c, c->keys, rec, (total + 4) * 32))
		return 1;
	c->n_v3_keys = n3;
	c->n_q_keys = nq;
	c->records_dirty = true;
	c->entries_dirty = true; // prepare_tables rebuilds the interval records
	CK(cudaStreamSynchronize(c->stream));
	invalidate_graph(c);
	return 0;
}

int hgb_set_tracks(hgb_ctx *c, uint32_t n, const hgb_track *tracks, uint32_t n_play) {
	if (!c)
		return fail("null context");
	join_post(c);
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
	for (uint32_t i = 0; i < n; ++i) {
		const uint64_t pool = tracks[i].kind == 3 ? c->n_q_keys : c->n_v3_keys;
		if (uint64_t(tracks[i].key_off) + tracks[i].key_count > pool)
			return fail("hgb_set_tracks: track %u reads keys [%u,%u) outside its pool of %llu (call hgb_set_anim_keys first)", i, tracks[i].key_off,
				tracks[i].key_off + tracks[i].key_count, (unsigned long long)pool);
	}
	{ // interval records are built per key range: two tracks may share a range (instanced animation) but ranges must not partially overlap
		std::vector<std::pair<uint64_t, uint32_t>> rng(n);
		for (uint32_t i = 0; i < n; ++i)
			rng[i] = {(uint64_t(tracks[i].kind == 3) << 32) | tracks[i].key_off, tracks[i].key_count};
		std::sort(rng.begin(), rng.end());
		for (uint32_t i = 1; i < n; ++i)
			if (rng[i] != rng[i - 1] && rng[i - 1].first + rng[i - 1].second > rng[i].first && (rng[i].first >> 32) == (rng[i - 1].first >> 32) &&
				rng[i - 1].second && rng[i].second)
				return fail("hgb_set_tracks: key ranges [%u,+%u) and [%u,+%u) overlap without being identical", uint32_t(rng[i - 1].first), rng[i - 1].second,
					uint32_t(rng[i].first), rng[i].second);
	}
	c->records_dirty = true;
	CK(c->play_t_d.reserve(std::max<size_t>(n_play, 1) * 8));
	CK(cudaStreamSynchronize(c->stream)); // no clock copy in flight reads the ring that is about to be replaced
	CK(c->play_t_h.reserve(std::max<size_t>(n_play, 1) * 8 * kClockRing));
	for (bool &u : c->clock_used)
		u = false;
	CK(c->cursor.reserve(std::max<size_t>(n, 1) * 4));
	CK(cudaMemsetAsync(c->cursor.p, 0, c->cursor.cap, c->stream));
	CK(c->tstate.reserve(std::max<size_t>(n, 1) * 16));
	CK(cudaMemsetAsync(c->tstate.p, 0, c->tstate.cap, c->stream));
	c->entries_dirty = true; // target positions are resolved with the entries
	invalidate_graph(c);
	return 0;
}

int hgb_set_enable_tracks(hgb_ctx *c, uint32_t n_tracks, const hgb_track *tracks, uint32_t n_keys, const int64_t *key_t, const uint8_t *key_v) {
	if (!c)
		return fail("null context");
	join_post(c);
	for (uint32_t i = 0; i < n_tracks; ++i) {
		if (tracks[i].target_slot >= c->n_slots)
			return fail("hgb_set_enable_tracks: track %u targets slot %u outside the topology", i, tracks[i].target_slot);
		if (tracks[i].play >= c->n_play)
			return fail("hgb_set_enable_tracks: track %u uses play slot %u of %u (hgb_set_tracks declares the play slots)", i, tracks[i].play, c->n_play);
		if (uint64_t(tracks[i].key_off) + tracks[i].key_count > n_keys)
			return fail("hgb_set_enable_tracks: track %u reads keys outside the pool", i);
	}
	if (fold_enable_state(c))
		return 1;
	CK(cudaSetDevice(c->device));
	c->en_tracks_h.assign(tracks, tracks + n_tracks);
	c->en_target_slot_h.clear();
	for (uint32_t i = 0; i < n_tracks; ++i)
		c->en_target_slot_h.push_back(tracks[i].target_slot);
	std::sort(c->en_target_slot_h.begin(), c->en_target_slot_h.end());
	c->en_target_slot_h.erase(std::unique(c->en_target_slot_h.begin(), c->en_target_slot_h.end()), c->en_target_slot_h.end());
	c->en_disabled_h.assign(c->en_target_slot_h.size(), 0);
	c->en_known_h.assign(c->en_target_slot_h.size(), 0);
	std::vector<uint4> keys(n_keys);
	for (uint32_t i = 0; i < n_keys; ++i)
		keys[i] = make_uint4(uint32_t(uint64_t(key_t[i])), uint32_t(uint64_t(key_t[i]) >> 32), key_v[i] ? 1u : 0u, 0u);
	if (upload(c, c->en_keys_d, keys.data(), size_t(n_keys) * 16))
		return 1;
	CK(cudaStreamSynchronize(c->stream));
	c->n_en_keys = n_keys;
	c->en_resolved = false;
	c->entries_dirty = true; // target positions are resolved with the entries
	invalidate_graph(c);
	return 0;
}

int hgb_set_enable_state(hgb_ctx *c, uint32_t n, const uint32_t *slot, const uint8_t *disabled) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (fold_enable_state(c)) // what the tracks did so far comes first, this explicit write lands after it
		return 1;
	for (uint32_t i = 0; i < n; ++i) {
		const auto it = std::lower_bound(c->en_target_slot_h.begin(), c->en_target_slot_h.end(), slot[i]);
		if (it == c->en_target_slot_h.end() || *it != slot[i])
			return fail("hgb_set_enable_state: slot %u is not the target of an enable track", slot[i]);
		const size_t k = it - c->en_target_slot_h.begin();
		c->en_disabled_h[k] = disabled[i] ? 1 : 0;
		c->en_known_h[k] = 1;
		for (int kind = 0; kind < 2; ++kind) {
			EntryTable &t = c->ent[kind];
			for (uint32_t e = 0; e < t.n; ++e)
				if (t.mtx_slot[e] == slot[i])
					t.flags[e] = uint8_t(disabled[i] ? (t.flags[e] | HGB_ENTRY_DISABLED) : (t.flags[e] & ~HGB_ENTRY_DISABLED));
		}
	}
	c->entries_dirty = true;
	invalidate_graph(c);
	return 0;
}

int hgb_get_enable_state(hgb_ctx *c, uint32_t capacity, uint32_t *out_slot, uint8_t *out_disabled, uint32_t *out_count) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (prepare_tables(c)) // tracks set but never resolved: resolve so that the device bits can be read
		return 1;
	c->en_state_dirty = c->en_state_dirty || !c->en_target_slot_h.empty();
	if (fold_enable_state(c))
		return 1;
	const uint32_t n = uint32_t(c->en_target_slot_h.size());
	if (out_count)
		*out_count = n;
	if (n > capacity)
		return fail("hgb_get_enable_state: %u tracked nodes but capacity %u", n, capacity);
	for (uint32_t i = 0; i < n; ++i)
		out_slot[i] = c->en_target_slot_h[i], out_disabled[i] = c->en_disabled_h[i];
	return 0;
}

int hgb_set_views(hgb_ctx *c, uint32_t n_views, const float *frustum24) {
	if (!c)
		return fail("null context");
	join_post(c);
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
		float condx[kMaxViews * 48];
		expand_planes(cond, n_views * 6, condx);
		CK(cudaMemcpyAsync(c->view_planes.p, cond, size_t(n_views) * 96, cudaMemcpyHostToDevice, c->stream));
		CK(cudaMemcpyAsync(c->view_planes_x.p, condx, size_t(n_views) * 192, cudaMemcpyHostToDevice, c->stream));
	}
	return 0;
}

int hgb_set_instance_views(hgb_ctx *c, uint32_t n_inst, const float *frustum24, const uint32_t *inst_entry_start) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (n_inst > 65535)
		return fail("hgb_set_instance_views: at most 65535 instances per context (uint16 view id)");
	CK(cudaSetDevice(c->device));
	const bool shape_change = !c->instance_mode || c->n_inst != n_inst;
	std::vector<float> cond(size_t(n_inst) * 24);
	precondition_planes(frustum24, n_inst * 6, cond.data());
	std::vector<float> condx(size_t(n_inst) * 48);
	expand_planes(cond.data(), n_inst * 6, condx.data());
	if (upload(c, c->inst_planes, cond.data(), size_t(n_inst) * 96) || upload(c, c->inst_planes_x, condx.data(), size_t(n_inst) * 192))
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
	++c->rebuilds;
	join_post(c); // the tables are about to be rebuilt
	if (fold_enable_state(c))
		return 1;
	if (c->n_bound) { // give the track-driven triples their values back before rows are rewritten and tracks re-bound
		k_bind_lanes<<<grid_for(c->n_bound), 256, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->anim_val.as<float4>(), c->bind_pos_d.as<uint32_t>(),
			c->bind_group_d.as<uint8_t>(), c->n_bound, 0);
		++c->launches;
		CK(cudaStreamSynchronize(c->stream));
		c->n_bound = 0;
	}
	const uint32_t n_pos_all = c->n_slots;
	std::vector<uint32_t> ent_off(size_t(n_pos_all) + 2, 0);
	for (int k = 0; k < 2; ++k) {
		EntryTable &t = c->ent[k];
		if (c->instance_mode && t.n && t.view.size() != t.n)
			return fail("instance views are set but entry table %d has no per-entry view ids", k);
		if (t.n_dev < t.n + c->append_slack)
			t.n_dev = t.n + c->append_slack;
		std::vector<uint32_t> mtx_pos(t.n_dev, 0u), meta(t.n_dev, 1u << 24); // free entries behind the table: disabled
		for (uint32_t e = 0; e < t.n; ++e) {
			if (t.mtx_slot[e] >= c->n_slots)
				return fail("display entry %u (kind %d) references slot %u outside the topology", e, k, t.mtx_slot[e]);
			if (t.bounds_idx[e] >= c->n_bounds)
				return fail("display entry %u (kind %d) references bounds %u outside the table", e, k, t.bounds_idx[e]);
			mtx_pos[e] = c->pos_of_slot[t.mtx_slot[e]];
			if (!pos_is_live(c, mtx_pos[e])) // its node was removed (hgb_remove_nodes) or never existed: the entry stays, disabled
				t.flags[e] |= HGB_ENTRY_DISABLED;
			else if (!c->tomb.empty()) { // ... also when the slot has been handed out again since
				const auto it = c->tomb.find(t.mtx_slot[e]);
				if (it != c->tomb.end() && e < it->second[k])
					t.flags[e] |= HGB_ENTRY_DISABLED;
			}
			meta[e] = t.bounds_idx[e] | (uint32_t(t.flags[e] & HGB_ENTRY_DISABLED ? 1 : 0) << 24);
			++ent_off[mtx_pos[e] + 1];
		}
		if (upload(c, t.mtx_pos, mtx_pos.data(), size_t(t.n_dev) * 4) || upload(c, t.meta, meta.data(), size_t(t.n_dev) * 4))
			return 1;
		if (!t.view.empty() && upload(c, t.view_d, t.view.data(), size_t(t.n) * 2))
			return 1;
		t.n_tiles = (t.n_dev + kCompactTile - 1) / kCompactTile;
		const size_t vis_bytes = size_t(t.n_tiles + 1) * kCompactTile;
		for (int par = 0; par < 2; ++par) {
			CK(t.vis[par].reserve(vis_bytes));
			CK(cudaMemsetAsync(t.vis[par].p, 0, t.vis[par].cap, c->stream));
		}
		const int n_out = c->instance_mode ? 1 : std::max(c->views.n_views, 1);
		for (int v = 0; v < n_out; ++v)
			CK(t.out[v].reserve(std::max<size_t>(t.n_dev, 1) * 4));
		CK(t.counts.reserve(kMaxViews * 4));
		CK(cudaMemsetAsync(t.counts.p, 0, t.counts.cap, c->stream));
		CK(t.ticket.reserve(kMaxViews * 4));
		CK(t.status.reserve(size_t(std::max(t.n_tiles, 1u)) * kMaxViews * 4));
		CK(cudaStreamSynchronize(c->stream));
	}
	// first entry of every position goes into tile rows 10/11; the others into the overflow CSR (ent_off counts them +1 so far)
	std::vector<uint32_t> row_meta(n_pos_all, 0), row_eidx(n_pos_all, 0);
	std::vector<uint16_t> pos_view(c->instance_mode ? n_pos_all : 0, 0);
	for (uint32_t p = 0; p < n_pos_all; ++p)
		if (ent_off[p + 1] > 0)
			--ent_off[p + 1]; // the first entry is not in the CSR
	for (size_t i = 0; i + 1 < ent_off.size(); ++i)
		ent_off[i + 1] += ent_off[i];
	const uint32_t total = ent_off[n_pos_all];
	std::vector<uint2> pent(total);
	{
		std::vector<uint32_t> fill(ent_off.begin(), ent_off.end() - 1);
		for (int k = 0; k < 2; ++k) {
			const EntryTable &t = c->ent[k];
			for (uint32_t e = 0; e < t.n; ++e) {
				const uint32_t p = c->pos_of_slot[t.mtx_slot[e]];
				const uint32_t meta = t.bounds_idx[e] | (t.flags[e] & HGB_ENTRY_DISABLED ? kEntDisabled : 0u);
				if (!(row_meta[p] & kEntHas)) {
					row_meta[p] = meta | kEntHas | (k ? kEntKindTransparent : 0u);
					row_eidx[p] = e;
					if (c->instance_mode)
						pos_view[p] = t.view[e];
				} else {
					row_meta[p] |= kEntMore;
					pent[fill[p]++] = make_uint2(e | (k ? kEntKindTransparent : 0u), meta);
					if (c->instance_mode && pos_view[p] != t.view[e])
						return fail("display entries of transform slot %u belong to different instances (%u and %u)", t.mtx_slot[e], pos_view[p], t.view[e]);
				}
			}
		}
	}
	// enable-tracked nodes that own no display entry keep their state in the meta row alone
	for (size_t i = 0; i < c->en_target_slot_h.size(); ++i) {
		if (c->en_target_slot_h[i] >= c->n_slots)
			return fail("enable track targets slot %u outside the topology", c->en_target_slot_h[i]);
		const uint32_t p = c->pos_of_slot[c->en_target_slot_h[i]];
		if (p < n_pos_all && !(row_meta[p] & kEntHas) && c->en_known_h[i] && c->en_disabled_h[i])
			row_meta[p] |= kEntDisabled;
	}
	if (upload(c, c->ent_off, ent_off.data(), (size_t(n_pos_all) + 1) * 4) || upload(c, c->pent, pent.data(), size_t(total) * 8))
		return 1;
	if (c->append_slack)
		for (uint32_t p = 0; p < c->cap_slots; ++p)
			c->has_entry_h[p] = p < n_pos_all && (row_meta[p] & kEntHas) ? 1 : 0;
	if (c->instance_mode && upload(c, c->pos_view, pos_view.data(), size_t(n_pos_all) * 2))
		return 1;
	if (n_pos_all) {
		DBuf rows;
		if (upload(c, rows, row_meta.data(), size_t(n_pos_all) * 4))
			return 1;
		k_fill_rows<<<grid_for(n_pos_all), 256, 0, c->stream>>>(c->tiles.as<uint32_t>(), kRowEntMeta, rows.as<uint32_t>(), n_pos_all);
		CK(cudaStreamSynchronize(c->stream));
		if (upload(c, rows, row_eidx.data(), size_t(n_pos_all) * 4))
			return 1;
		k_fill_rows<<<grid_for(n_pos_all), 256, 0, c->stream>>>(c->tiles.as<uint32_t>(), kRowEntIdx, rows.as<uint32_t>(), n_pos_all);
		CK(cudaStreamSynchronize(c->stream));
		rows.release();
		c->launches += 2;
	}

	{ // enable tracks: CSR by target node, evaluation order kept inside a target
		const uint32_t nt = uint32_t(c->en_target_slot_h.size());
		std::vector<uint32_t> tpos(nt), off(size_t(nt) + 1, 0);
		for (uint32_t i = 0; i < nt; ++i)
			tpos[i] = c->pos_of_slot[c->en_target_slot_h[i]];
		std::vector<uint32_t> tgt_of(c->en_tracks_h.size());
		for (size_t k = 0; k < c->en_tracks_h.size(); ++k) {
			tgt_of[k] = uint32_t(std::lower_bound(c->en_target_slot_h.begin(), c->en_target_slot_h.end(), c->en_tracks_h[k].target_slot) - c->en_target_slot_h.begin());
			++off[tgt_of[k] + 1];
		}
		for (uint32_t i = 0; i < nt; ++i)
			off[i + 1] += off[i];
		std::vector<uint4> trk(c->en_tracks_h.size());
		std::vector<uint32_t> fill(off.begin(), off.end() - 1);
		for (size_t k = 0; k < c->en_tracks_h.size(); ++k)
			trk[fill[tgt_of[k]]++] = make_uint4(c->en_tracks_h[k].key_off, c->en_tracks_h[k].key_count, c->en_tracks_h[k].play, 0u);
		if (upload(c, c->en_target_pos, tpos.data(), size_t(nt) * 4) || upload(c, c->en_off, off.data(), off.size() * 4) ||
			upload(c, c->en_tracks_d, trk.data(), trk.size() * 16))
			return 1;
		CK(cudaStreamSynchronize(c->stream));
		c->en_resolved = true;
	}

	// bones and tracks: slot -> position
	std::vector<uint32_t> bone_pos(c->n_bones);
	for (uint32_t i = 0; i < c->n_bones; ++i) {
		if (c->bone_mtx_slot[i] >= c->n_slots)
			return fail("skin record %u references slot %u outside the topology", i, c->bone_mtx_slot[i]);
		bone_pos[i] = c->pos_of_slot[c->bone_mtx_slot[i]];
	}
	if (upload(c, c->bone_mtx_pos, bone_pos.data(), size_t(c->n_bones) * 4))
		return 1;
	// tracks: every driven (position, pos|rot|scale) triple gets one anim_val slot, ordered by position; tracks are sorted by slot
	// inside their write round so that consecutive threads write consecutive float4s
	std::vector<Track> tr(c->tracks_h.size());
	std::vector<uint64_t> keys_pg(tr.size());
	for (size_t i = 0; i < tr.size(); ++i) {
		const hgb_track &h = c->tracks_h[i];
		if (h.target_slot >= c->n_slots)
			return fail("track %zu targets slot %u outside the topology", i, h.target_slot);
		keys_pg[i] = (uint64_t(c->pos_of_slot[h.target_slot]) << 2) | (h.kind == 3 ? 1u : h.kind);
	}
	std::vector<uint64_t> uniq(keys_pg);
	std::sort(uniq.begin(), uniq.end());
	uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
	for (size_t i = 0; i < tr.size(); ++i) {
		const hgb_track &h = c->tracks_h[i];
		const uint32_t slot = uint32_t(std::lower_bound(uniq.begin(), uniq.end(), keys_pg[i]) - uniq.begin());
		tr[i] = Track{slot, h.key_off + (h.kind == 3 ? c->n_v3_keys : 0u), h.key_count, h.play, h.kind, 1}; // key_off: compact-pool key for now
	}
	for (size_t r = 0; r + 1 < c->round_off.size(); ++r)
		std::sort(tr.begin() + c->round_off[r], tr.begin() + c->round_off[r + 1], [](const Track &a, const Track &b) { return a.out_slot < b.out_slot; });
	// interval records: neighbour tracks of a round are paired and share their 128-byte lines (see k_build_key_records)
	std::vector<TrackBuild> tb(tr.size());
	uint64_t rec_units = 0;
	for (size_t r = 0; r + 1 < c->round_off.size(); ++r)
		for (size_t i = c->round_off[r]; i < c->round_off[r + 1];) {
			const bool pair = i + 1 < c->round_off[r + 1];
			const uint32_t n0 = tr[i].key_count, n1 = pair ? tr[i + 1].key_count : 0u;
			for (int h = 0; h < (pair ? 2 : 1); ++h) {
				Track &t_ = tr[i + h];
				tb[i + h] = TrackBuild{t_.key_off, uint32_t(rec_units) + uint32_t(h), t_.key_count, t_.kind, uint8_t(pair ? 2 : 1), {0, 0}};
				t_.key_off = uint32_t(rec_units) + uint32_t(h);
				t_.stride = uint8_t(pair ? 2 : 1);
			}
			rec_units += pair ? 2ull * std::max(n0, n1) : n0;
			i += pair ? 2 : 1;
		}
	if (rec_units >= (1ull << 32))
		return fail("the interval records of the bound tracks exceed 2^32 x 64 bytes");
	if (upload(c, c->tracks_d, tr.data(), tr.size() * sizeof(Track)))
		return 1;
	// chunk boundaries of the overlapped sampler: target position of every round-0 track (ascending: tracks are sorted by slot = position)
	c->trk_pos_h.clear();
	if (c->round_off.size() == 2 && !tr.empty()) {
		c->trk_pos_h.resize(tr.size());
		for (size_t i = 0; i < tr.size(); ++i)
			c->trk_pos_h[i] = uint32_t(uniq[tr[i].out_slot] >> 2);
	}
	if (!tr.empty()) { // 64-byte interval records from the compact key pool (device side, at every re-bind: the layout follows the track order)
		DBuf tb_d;
		if (upload(c, tb_d, tb.data(), tb.size() * sizeof(TrackBuild)))
			return 1;
		CK(c->key_rec.reserve((rec_units + 2) * 64));
		k_build_key_records<<<grid_for(uint32_t(tr.size())), 256, 0, c->stream>>>(tb_d.as<TrackBuild>(), uint32_t(tr.size()), c->keys.as<uint4>(), c->key_rec.as<uint4>());
		++c->launches;
		CK(cudaGetLastError());
		CK(cudaStreamSynchronize(c->stream));
		tb_d.release();
		c->records_dirty = false;
	}
	const uint32_t n_bind = uint32_t(uniq.size());
	if (n_bind) {
		std::vector<uint32_t> bpos(n_bind);
		std::vector<uint8_t> bgrp(n_bind);
		for (uint32_t i = 0; i < n_bind; ++i)
			bpos[i] = uint32_t(uniq[i] >> 2), bgrp[i] = uint8_t(uniq[i] & 3);
		if (upload(c, c->bind_pos_d, bpos.data(), size_t(n_bind) * 4) || upload(c, c->bind_group_d, bgrp.data(), n_bind))
			return 1;
		CK(c->anim_val.reserve(size_t(n_bind) * 16));
		CK(cudaMemsetAsync(c->cursor.p, 0, c->cursor.cap, c->stream));
		CK(cudaMemsetAsync(c->tstate.p, 0, c->tstate.cap, c->stream));
		k_bind_lanes<<<grid_for(n_bind), 256, 0, c->stream>>>(c->tiles.as<uint32_t>(), c->anim_val.as<float4>(), c->bind_pos_d.as<uint32_t>(),
			c->bind_group_d.as<uint8_t>(), n_bind, 1);
		++c->launches;
		CK(cudaStreamSynchronize(c->stream)); // bpos/bgrp go out of scope
		c->n_bound = n_bind;
	}
	CK(cudaStreamSynchronize(c->stream));
	c->entries_dirty = false;
	invalidate_graph(c);
	return 0;
}
