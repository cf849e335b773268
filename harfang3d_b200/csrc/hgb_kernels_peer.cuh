// Multi-GPU result gather over NVLink / NVSwitch peer memory. Scene instances shard across GPUs with no data-path
// exchange; the only exchange of the path is the gather of its RESULT (visible counts and lists) to one rank. Instead of a host-sized
// collective after the frame (counts -> host -> ncclSend/ncclRecv), the producing kernel writes over the wire: k_compact stores every
// survivor both into the local list and, globalised, into this rank's region of the destination GPU's inbox (peer stores through the
// NVSwitch fabric), a one-thread kernel then publishes count + step flag, and the destination concatenates the regions once all flags
// of the step are up. Everything is enqueued on the contexts' streams (and captured in the frame graph): no host synchronisation, no
// host-side sizes. Flow control: two inbox halves used alternately (the parity of the gather's step counter); a sender may reuse a half
// only after the destination acknowledged the step that used it (ack word in the sender's mailbox, written by the destination).
// Spins are bounded (about two seconds of SM clocks): a missing peer sets an error word instead of hanging the device.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hgb {

constexpr int kMaxPeers = 16;
constexpr uint32_t kPeerHdrWords = 32; // one 128-byte line per (half, rank): [0] step flag, [1] visible count
constexpr long long kPeerSpinClocks = 4LL << 30;

// mailbox words (local memory of every rank): [0] ack (written by the destination), [1] step counter, [2] error code.
// The inbox HALF of a step is its parity (step & 1), selected on the device from the step word: whatever else the host does between
// two gathers (frames without a gather, other world-buffer parities), flow control and buffer choice count the same thing.
__device__ __forceinline__ bool peer_spin(const volatile uint32_t *w, uint32_t at_least, uint32_t *err, uint32_t code) {
	const long long t0 = clock64();
	while (*w < at_least) {
		__nanosleep(100);
		if (clock64() - t0 > kPeerSpinClocks) {
			*err = code;
			return false;
		}
	}
	return true;
}

// first kernel of a gather (ahead of the compaction that stores into the inbox): count the step, wait until its half is free again
__global__ void k_peer_begin(uint32_t *mailbox) {
	const uint32_t step = mailbox[1] + 1;
	mailbox[1] = step;
	if (step > 2)
		peer_spin(mailbox, step - 2, mailbox + 2, 1u);
}

// aux words (per-instance ranks) of this rank into its region of the step's half
__global__ void __launch_bounds__(256) k_peer_copy(uint32_t *__restrict__ inbox, uint32_t half_words, uint32_t off, const uint32_t *__restrict__ src, uint32_t n,
	const uint32_t *__restrict__ mailbox) {
	uint32_t *dst = inbox + size_t(mailbox[1] & 1u) * half_words + off;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
		dst[i] = src[i];
}

// after the compaction (and the aux copy) of this step: their peer stores are performed (kernel boundary + fence), then the flag goes up
__global__ void k_peer_publish(uint32_t *inbox, uint32_t half_words, uint32_t hdr_off, const uint32_t *count, const uint32_t *mailbox) {
	const uint32_t step = mailbox[1];
	uint32_t *hdr = inbox + size_t(step & 1u) * half_words + hdr_off;
	__threadfence_system();
	hdr[1] = count ? *count : 0u; // a rank without display entries still takes part, with an empty list
	__threadfence_system();
	*reinterpret_cast<volatile uint32_t *>(hdr) = step;
}

struct PeerConcat {
	const uint32_t *inbox;        // base of the inbox (both halves)
	uint32_t half_words;
	uint32_t world;
	uint32_t list_off[kMaxPeers]; // word offsets inside a half
	uint32_t aux_off[kMaxPeers];
	uint32_t aux_n[kMaxPeers], aux_base[kMaxPeers + 1]; // aux words per rank and their prefix
	uint32_t *out_list;           // concatenated visible entries (global indices), rank order
	uint32_t *rank_off;           // world + 1 exclusive offsets
	uint32_t *aux_out;            // aux of all ranks, each value + rank_off[r] (instance ranks -> global offsets), then the total
	const uint32_t *mailbox;
};

// destination: wait for the flags of this step
__global__ void k_peer_wait(const uint32_t *inbox, uint32_t half_words, uint32_t world, uint32_t *mailbox) {
	if (threadIdx.x < world) {
		const uint32_t step = mailbox[1];
		peer_spin(inbox + size_t(step & 1u) * half_words + threadIdx.x * kPeerHdrWords, step, mailbox + 2, 2u + threadIdx.x);
	}
}

__global__ void __launch_bounds__(256) k_peer_concat(PeerConcat p) {
	__shared__ uint32_t off[kMaxPeers + 1];
	if (p.mailbox[2] != 0) // a peer timed out: the regions are not this step's — keep the previous result, the error word tells
		return;
	const uint32_t *half = p.inbox + size_t(p.mailbox[1] & 1u) * p.half_words;
	if (threadIdx.x == 0) {
		uint32_t acc = 0;
		for (uint32_t r = 0; r < p.world; ++r) {
			off[r] = acc;
			acc += half[r * kPeerHdrWords + 1];
		}
		off[p.world] = acc;
	}
	__syncthreads();
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
	for (uint32_t r = 0; r < p.world; ++r) {
		const uint32_t *src = half + p.list_off[r];
		uint32_t *dst = p.out_list + off[r];
		const uint32_t n = off[r + 1] - off[r];
		for (uint32_t i = tid; i < n; i += stride)
			dst[i] = src[i];
		const uint32_t *ax = half + p.aux_off[r];
		for (uint32_t i = tid; i < p.aux_n[r]; i += stride)
			p.aux_out[p.aux_base[r] + i] = ax[i] + off[r];
	}
	if (tid <= p.world)
		p.rank_off[tid] = off[tid];
	if (tid == 0)
		p.aux_out[p.aux_base[p.world]] = off[p.world];
}

struct PeerAck {
	uint32_t *mailbox[kMaxPeers]; // ack word of every rank (peer pointers; own rank: local)
	uint32_t world;
};
__global__ void k_peer_ack(PeerAck a, const uint32_t *own_mailbox) {
	if (threadIdx.x < a.world && own_mailbox[2] == 0) { // after a timeout nothing is acknowledged: the senders time out too instead of overwriting
		__threadfence_system();
		*reinterpret_cast<volatile uint32_t *>(a.mailbox[threadIdx.x]) = own_mailbox[1];
	}
}

} // namespace hgb
