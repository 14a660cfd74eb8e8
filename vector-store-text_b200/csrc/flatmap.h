// flatmap.h — u64 key -> u32 slot, open addressing with linear probing and backward-shift deletion.
// The host half of add / replace / remove (src/index/usearch.rs:181-196,214-225,240) touches this map once per vector; at
// 2 M vectors/s of GPU build a node-based std::unordered_map (~150 ns and one allocation per insert) was a fifth of the wall
// clock, so the map is a flat table: one cache line per probe, no allocation per key.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace vsann {

class FlatMap {
public:
    static constexpr uint32_t NO_SLOT = 0xFFFFFFFFu;
    FlatMap() { rehash(1024); }

    size_t size() const { return n_; }

    void reserve(size_t want) { if (want * 10 > cap_ * 6) { size_t c = cap_; while (want * 10 > c * 6) c <<= 1; rehash(c); } }

    // Returns the previous slot of `key` (NO_SLOT if it was absent) and maps it to `slot`.
    uint32_t put(uint64_t key, uint32_t slot) {
        if ((n_ + 1) * 10 > cap_ * 6) rehash(cap_ << 1);
        size_t i = hash(key) & (cap_ - 1);
        for (;;) {
            if (val_[i] == NO_SLOT) { key_[i] = key; val_[i] = slot; ++n_; return NO_SLOT; }
            if (key_[i] == key) { const uint32_t old = val_[i]; val_[i] = slot; return old; }
            i = (i + 1) & (cap_ - 1);
        }
    }

    uint32_t get(uint64_t key) const {
        size_t i = hash(key) & (cap_ - 1);
        for (;;) {
            if (val_[i] == NO_SLOT) return NO_SLOT;
            if (key_[i] == key) return val_[i];
            i = (i + 1) & (cap_ - 1);
        }
    }

    // Removes `key`; returns its slot or NO_SLOT.
    uint32_t erase(uint64_t key) {
        size_t i = hash(key) & (cap_ - 1);
        for (;;) {
            if (val_[i] == NO_SLOT) return NO_SLOT;
            if (key_[i] == key) break;
            i = (i + 1) & (cap_ - 1);
        }
        const uint32_t old = val_[i];
        size_t j = i;                                  // backward shift: close the gap so that probes never cross an empty cell
        for (;;) {
            j = (j + 1) & (cap_ - 1);
            if (val_[j] == NO_SLOT) break;
            const size_t home = hash(key_[j]) & (cap_ - 1);
            const bool between = (i <= j) ? (i < home && home <= j) : (i < home || home <= j);
            if (!between) { key_[i] = key_[j]; val_[i] = val_[j]; i = j; }
        }
        val_[i] = NO_SLOT; --n_;
        return old;
    }

    void clear() { std::fill(val_.begin(), val_.end(), NO_SLOT); n_ = 0; }

    template <class F> void for_each(F f) const { for (size_t i = 0; i < cap_; ++i) if (val_[i] != NO_SLOT) f(key_[i], val_[i]); }
    template <class F> void update_each(F f) { for (size_t i = 0; i < cap_; ++i) if (val_[i] != NO_SLOT) val_[i] = f(key_[i], val_[i]); }

private:
    static size_t hash(uint64_t k) { k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33; return (size_t)k; }
    void rehash(size_t c) {
        std::vector<uint64_t> ok; std::vector<uint32_t> ov;
        ok.swap(key_); ov.swap(val_);
        key_.assign(c, 0); val_.assign(c, NO_SLOT); cap_ = c; n_ = 0;
        for (size_t i = 0; i < ov.size(); ++i) if (ov[i] != NO_SLOT) put(ok[i], ov[i]);
    }
    std::vector<uint64_t> key_; std::vector<uint32_t> val_; size_t cap_ = 0, n_ = 0;
};

}  // namespace vsann
