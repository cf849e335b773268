// Host side of the peer-memory result gather (hgb_kernels_peer.cuh). Included by hgb_ctx.cu.


extern "C" {

int hgb_peer_init(hgb_ctx *c, uint32_t rank, uint32_t world, uint32_t dst, const uint32_t *entries_per_rank, const uint32_t *aux_words_per_rank,
	hgb_peer_handle *inbox_out, hgb_peer_handle *mailbox_out) {
	if (!c)
		return fail("null context");
	join_post(c);
	if (world == 0 || world > uint32_t(kMaxPeers) || rank >= world || dst >= world)
		return fail("hgb_peer_init: bad rank / world / dst (at most %d ranks)", kMaxPeers);
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	peer_release(c); // a second init replaces the first
	invalidate_graph(c);
	Peer &p = c->peer;
	p.rank = rank, p.world = world, p.dst = dst;
	// one inbox half: world header lines, then every rank's list region (capacity = its entry count) and aux region, 128-byte aligned
	size_t w = size_t(world) * kPeerHdrWords;
	uint64_t base = 0;
	for (uint32_t r = 0; r < world; ++r) {
		p.cap[r] = entries_per_rank[r];
		p.aux_n[r] = aux_words_per_rank ? aux_words_per_rank[r] : 0;
		p.entry_base[r] = uint32_t(base);
		base += p.cap[r];
		p.list_off[r] = uint32_t(w);
		w += (size_t(p.cap[r]) + 31) & ~size_t(31);
		p.aux_off[r] = uint32_t(w);
		w += (size_t(p.aux_n[r]) + 31) & ~size_t(31);
	}
	if (base >= (1ull << 32) || w >= (1ull << 30))
		return fail("hgb_peer_init: the gathered entry table exceeds 32-bit indices");
	p.half_words = w;
	p.aux_base[0] = 0;
	for (uint32_t r = 0; r < world; ++r)
		p.aux_base[r + 1] = p.aux_base[r] + p.aux_n[r];
	CK(cudaMalloc(&p.mailbox, 128));
	CK(cudaMemset(p.mailbox, 0, 128));
	cudaIpcMemHandle_t h;
	CK(cudaIpcGetMemHandle(&h, p.mailbox));
	static_assert(sizeof(h) == sizeof(hgb_peer_handle), "cudaIpcMemHandle_t is 64 bytes");
	memcpy(mailbox_out, &h, sizeof(h));
	memset(inbox_out, 0, sizeof(*inbox_out));
	if (rank == dst) {
		CK(cudaMalloc(&p.inbox_local, w * 2 * 4));
		CK(cudaMemset(p.inbox_local, 0, w * 2 * 4));
		CK(cudaIpcGetMemHandle(&h, p.inbox_local));
		memcpy(inbox_out, &h, sizeof(h));
		p.inbox = static_cast<uint32_t *>(p.inbox_local);
		CK(c->peer_out_list.reserve(std::max<uint64_t>(base, 1) * 4));
		CK(c->peer_rank_off.reserve((size_t(world) + 1) * 4));
		CK(c->peer_aux_out.reserve((size_t(p.aux_base[world]) + 1) * 4));
	}
	CK(cudaDeviceSynchronize());
	return 0;
}

int hgb_peer_connect(hgb_ctx *c, const hgb_peer_handle *inbox_of_dst, const hgb_peer_handle *mailboxes) {
	if (!c)
		return fail("null context");
	join_post(c);
	Peer &p = c->peer;
	if (!p.world || !p.mailbox)
		return fail("hgb_peer_connect: hgb_peer_init first");
	CK(cudaSetDevice(c->device));
	if (p.rank != p.dst) {
		cudaIpcMemHandle_t h;
		memcpy(&h, inbox_of_dst, sizeof(h));
		void *ptr = nullptr;
		CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
		p.inbox = static_cast<uint32_t *>(ptr);
		p.opened.push_back(ptr);
	} else {
		for (uint32_t r = 0; r < p.world; ++r) {
			if (r == p.rank) {
				p.ack_ptr[r] = static_cast<uint32_t *>(p.mailbox);
				continue;
			}
			cudaIpcMemHandle_t h;
			memcpy(&h, mailboxes + r, sizeof(h));
			void *ptr = nullptr;
			CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
			p.ack_ptr[r] = static_cast<uint32_t *>(ptr);
			p.opened.push_back(ptr);
		}
	}
	p.on = true;
	invalidate_graph(c);
	return 0;
}

int hgb_peer_close(hgb_ctx *c) {
	if (!c)
		return fail("null context");
	join_post(c);
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	peer_release(c);
	invalidate_graph(c);
	return 0;
}

int hgb_peer_status(hgb_ctx *c, uint32_t *out_step, uint32_t *out_error) {
	if (!c)
		return fail("null context");
	join_post(c);
	Peer &p = c->peer;
	if (!p.mailbox)
		return fail("hgb_peer_status: no peer gather on this context");
	CK(cudaSetDevice(c->device));
	uint32_t w[4] = {};
	CK(cudaMemcpyAsync(w, p.mailbox, 16, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	if (out_step)
		*out_step = w[1];
	if (out_error)
		*out_error = w[2];
	return 0;
}

int hgb_peer_result(hgb_ctx *c, uint32_t *out_rank_off, uint32_t *out_idx, uint32_t idx_capacity, uint32_t *out_aux, uint32_t aux_capacity) {
	if (!c)
		return fail("null context");
	c->result_read = true;
	join_post(c);
	Peer &p = c->peer;
	if (!p.on || p.rank != p.dst)
		return fail("hgb_peer_result: only the gather destination holds the result");
	CK(cudaSetDevice(c->device));
	std::vector<uint32_t> off(p.world + 1);
	uint32_t mw[4] = {};
	CK(cudaMemcpyAsync(mw, p.mailbox, 16, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaMemcpyAsync(off.data(), c->peer_rank_off.p, off.size() * 4, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	if (mw[2] != 0)
		return fail("hgb_peer_result: the gather failed at step %u (error word %u: %s); the result on the device is not this step's", mw[1], mw[2],
			mw[2] == 1 ? "a sender timed out waiting for an acknowledgement" : "the destination timed out waiting for a rank");
	if (out_rank_off)
		memcpy(out_rank_off, off.data(), off.size() * 4);
	if (out_idx) {
		if (off[p.world] > idx_capacity)
			return fail("hgb_peer_result: %u entries but capacity %u", off[p.world], idx_capacity);
		CK(cudaMemcpyAsync(out_idx, c->peer_out_list.p, size_t(off[p.world]) * 4, cudaMemcpyDeviceToHost, c->stream));
	}
	if (out_aux) {
		if (p.aux_base[p.world] + 1 > aux_capacity)
			return fail("hgb_peer_result: %u aux words but capacity %u", p.aux_base[p.world] + 1, aux_capacity);
		CK(cudaMemcpyAsync(out_aux, c->peer_aux_out.p, (size_t(p.aux_base[p.world]) + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
	}
	CK(cudaStreamSynchronize(c->stream));
	return 0;
}

} // extern "C"

static void peer_release(hgb_ctx *c) {
	Peer &p = c->peer;
	for (void *q : p.opened)
		cudaIpcCloseMemHandle(q);
	if (p.mailbox)
		cudaFree(p.mailbox);
	if (p.inbox_local)
		cudaFree(p.inbox_local);
	p = Peer{};
}

// ---- per frame ----
static void peer_enqueue_begin(hgb_ctx *c, uint64_t *n_k) {
	k_peer_begin<<<1, 1, 0, c->stream>>>(static_cast<uint32_t *>(c->peer.mailbox));
	++*n_k;
}

// after the compaction of the frame: aux (per-instance ranks), count + flag; on the destination: wait, concatenate, acknowledge.
// The inbox half is chosen on the device from the step counter (k_peer_begin), never from host state.
static int peer_enqueue_end(hgb_ctx *c, uint64_t *n_k) {
	Peer &p = c->peer;
	uint32_t *mail = static_cast<uint32_t *>(p.mailbox);
	const uint32_t half_words = uint32_t(p.half_words);
	if (p.aux_n[p.rank]) {
		if (!c->instance_mode || p.aux_n[p.rank] != c->n_inst)
			return fail("peer gather: %u aux words declared but the context has %u instances", p.aux_n[p.rank], c->instance_mode ? c->n_inst : 0);
		k_peer_copy<<<std::min<uint32_t>((p.aux_n[p.rank] + 255) / 256, 148u), 256, 0, c->stream>>>(p.inbox, half_words, p.aux_off[p.rank], c->inst_rank.as<uint32_t>(),
			p.aux_n[p.rank], mail);
		++*n_k;
	}
	k_peer_publish<<<1, 1, 0, c->stream>>>(p.inbox, half_words, p.rank * kPeerHdrWords, c->ent[0].n_dev ? c->ent[0].counts.as<uint32_t>() : nullptr, mail);
	++*n_k;
	if (p.rank == p.dst) {
		k_peer_wait<<<1, 32, 0, c->stream>>>(p.inbox, half_words, p.world, mail);
		PeerConcat pc{};
		pc.inbox = p.inbox;
		pc.half_words = half_words;
		pc.world = p.world;
		for (uint32_t r = 0; r < p.world; ++r)
			pc.list_off[r] = p.list_off[r], pc.aux_off[r] = p.aux_off[r], pc.aux_n[r] = p.aux_n[r];
		for (uint32_t r = 0; r <= p.world; ++r)
			pc.aux_base[r] = p.aux_base[r];
		pc.out_list = c->peer_out_list.as<uint32_t>();
		pc.rank_off = c->peer_rank_off.as<uint32_t>();
		pc.aux_out = c->peer_aux_out.as<uint32_t>();
		pc.mailbox = mail;
		k_peer_concat<<<148 * 4, 256, 0, c->stream>>>(pc);
		PeerAck a{};
		a.world = p.world;
		for (uint32_t r = 0; r < p.world; ++r)
			a.mailbox[r] = p.ack_ptr[r];
		k_peer_ack<<<1, 32, 0, c->stream>>>(a, mail);
		*n_k += 3;
	}
	CK(cudaGetLastError());
	return 0;
}
