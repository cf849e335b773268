// hg::Scene facade — index-stable slot containers with the reference's OBSERVABLE behaviour (harfang/foundation/vector_list.h:28-355,
// generational_vector_list.h:10-68): slots are handed out in ascending order, a removed slot becomes the next one handed out
// (LIFO free list), capacity grows as capacity*2+16, iteration visits used slots in index order, a generation counter per slot
// invalidates stale gen_refs. Slot indices appear in the outputs of the path (ModelDisplayList::mtx_idx, world-matrix indices),
// so this order is part of the contract. The implementation is independent: a used-flag vector and an explicit free chain.
#pragma once
#include <cstdint>
#include <utility>
#include <vector>

namespace hg {

struct gen_ref {
	uint32_t idx{0xffffffff}, gen{0xffffffff};
};
static const gen_ref invalid_gen_ref;
inline bool operator==(gen_ref a, gen_ref b) { return a.idx == b.idx && a.gen == b.gen; }
inline bool operator!=(gen_ref a, gen_ref b) { return !(a == b); }
inline bool operator<(gen_ref a, gen_ref b) { return a.gen == b.gen ? a.idx < b.idx : a.gen < b.gen; }

template <typename T> class vector_list {
public:
	static const uint32_t invalid_idx = 0xffffffff;

	void reserve(size_t count) {
		const size_t cap = items_.size();
		if (count <= cap)
			return;
		items_.resize(count);
		used_.resize(count, 0);
		next_free_.resize(count);
		for (size_t i = cap; i < count; ++i)
			next_free_[i] = uint32_t(i + 1); // fresh slots chain upwards; the old tail already points at `cap`
	}
	size_t size() const { return size_; }
	size_t capacity() const { return items_.size(); }
	void clear() {
		items_.clear();
		used_.clear();
		next_free_.clear();
		size_ = 0;
		free_ = 0;
	}

	T &operator[](size_t i) { return items_[i]; }
	const T &operator[](size_t i) const { return items_[i]; }

	bool is_used(uint32_t i) const { return i < used_.size() && used_[i]; }
	uint32_t first() const { return scan(0); }
	uint32_t next(uint32_t from) const { return scan(size_t(from) + 1); }

	uint32_t add(T v) {
		if (size_ == capacity())
			reserve(capacity() * 2 + 16);
		const uint32_t i = free_;
		free_ = next_free_[i];
		items_[i] = std::move(v);
		used_[i] = 1;
		++size_;
		return i;
	}
	uint32_t remove(uint32_t i) { // returns the next used index, like the reference
		const uint32_t n = next(i);
		items_[i] = T();
		used_[i] = 0;
		next_free_[i] = free_;
		free_ = i;
		--size_;
		return n;
	}

private:
	uint32_t scan(size_t i) const {
		for (; i < used_.size(); ++i)
			if (used_[i])
				return uint32_t(i);
		return invalid_idx;
	}
	std::vector<T> items_;
	std::vector<uint8_t> used_;
	std::vector<uint32_t> next_free_;
	size_t size_{0};
	uint32_t free_{0};
};

template <typename T> class generational_vector_list : public vector_list<T> {
public:
	gen_ref add_ref(T v) {
		const uint32_t idx = vector_list<T>::add(std::move(v));
		if (idx >= generations_.size())
			generations_.resize(size_t(idx) + 64, 0);
		return {idx, generations_[idx]};
	}
	void remove_ref(gen_ref ref) {
		if (is_valid(ref)) {
			++generations_[ref.idx];
			this->remove(ref.idx);
		}
	}
	gen_ref first_ref() const { return get_ref_or_invalid(this->first()); }
	gen_ref next_ref(gen_ref ref) const { return get_ref_or_invalid(this->next(ref.idx)); }
	gen_ref get_ref(uint32_t idx) const { return this->is_used(idx) && idx < generations_.size() ? gen_ref{idx, generations_[idx]} : invalid_gen_ref; }
	bool is_valid(gen_ref ref) const { return this->is_used(ref.idx) && ref.idx < generations_.size() && ref.gen == generations_[ref.idx]; }
	void clear() {
		vector_list<T>::clear();
		generations_.clear();
	}

private:
	gen_ref get_ref_or_invalid(uint32_t idx) const { return idx == 0xffffffff ? invalid_gen_ref : gen_ref{idx, generations_[idx]}; }
	std::vector<uint32_t> generations_;
};

} // namespace hg
