// rbeam.cuh — register-resident candidate lists for the warp beam (ef <= 32 * NL), striped across the warp: entry i sits
// in row i / 32 of lane i % 32. Two implementations with one interface (RList<NL> picks by width, measured on B200 at 1M x 128:
// NL = 1: sorted 0.506 ms vs unsorted 0.533 ms per 10k queries; NL = 2 equal; NL = 4 (ef_add = 128 build) unsorted 1.6 % faster,
// and its cost per operation no longer grows with NL):
//
// SortedList - kept sorted: popping the next node is one ballot per row, inserting a candidate one ballot + four shuffles
// per row. Cheap for one or two rows.
//
// UList - NOT kept sorted: a beam needs three things from it - the worst kept distance (the admission
// threshold), the nearest unexpanded entry (the next node to expand) and, once per query, the entries in ascending order.
// Distances are held as order-preserving 32-bit keys, so each of the first two is one pass over the lane's NL registers
// plus ONE warp reduction instruction (REDUX.MIN / REDUX.MAX, `__reduce_min_sync`): about NL + 6 instructions, against
// one ballot + four shuffles per row per inserted candidate for the sorted list this replaces (ncu on the 1M x 128 build,
// ef_add = 128, NL = 4: 121 k warp instructions per inserted vector, most of them list maintenance). Inserting overwrites
// the worst entry in place; nothing shifts.
// Same set semantics as SList (beam.cuh) and as the oracle's two-heap beam (oracle_vsann.cpp search_layer): the list holds
// the best `cap` candidates seen, ties ordered by (distance, id); the beam ends when every listed entry is expanded.
#pragma once
#include "beam.cuh"

namespace vsann {

constexpr uint32_t TAKEN = 0x40000000u;      // flag bit used while the list is read out in ascending order (slots < 2^30)
constexpr uint32_t ID_MASK = 0x3FFFFFFFu;
constexpr uint32_t KEY_NONE = 0xFFFFFFFFu;   // above the key of +inf (0xFF800000)

// float <-> u32 key with the same order (negative values included: IP and cosine distances may dip below zero)
__device__ __forceinline__ uint32_t fkey(float f) { const uint32_t b = __float_as_uint(f); return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u); }
__device__ __forceinline__ float fkey_inv(uint32_t k) { return __uint_as_float(k ^ ((k & 0x80000000u) ? 0x80000000u : 0xFFFFFFFFu)); }

template <int NL>
struct SortedList {
    float d[NL];
    uint32_t id[NL];
    int cap;
    float worst;   // distance of entry cap-1 (+inf while the list is not full), warp-uniform

    __device__ __forceinline__ void init(int cap_) {
        cap = cap_; worst = INFINITY;
#pragma unroll
        for (int j = 0; j < NL; ++j) { d[j] = INFINITY; id[j] = NONE; }
    }

    __device__ __forceinline__ void seed(float nd, uint32_t nid) {
        if (lane_id() == 0) { d[0] = nd; id[0] = nid; }
        refresh_worst();
    }

    __device__ __forceinline__ void refresh_worst() {
        const int r = (cap - 1) >> 5, l = (cap - 1) & 31;
        float w = INFINITY;
#pragma unroll
        for (int j = 0; j < NL; ++j) { const float v = __shfl_sync(0xffffffffu, d[j], l); if (j == r) w = v; }
        worst = w;
    }

    // nd / nid are warp-uniform. Drops duplicates and anything that does not beat `worst`.
    __device__ __forceinline__ void insert(float nd, uint32_t nid) {
        const unsigned lane = lane_id();
        bool dup = false;
#pragma unroll
        for (int j = 0; j < NL; ++j) dup |= ((id[j] & ~EXPANDED) == nid) && id[j] != NONE;
        if (__any_sync(0xffffffffu, dup)) return;
        bool shifting = false;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            int pos;
            if (shifting) pos = 0;
            else pos = __popc(__ballot_sync(0xffffffffu, pair_less(d[j], id[j] & ~EXPANDED, nd, nid)));
            const float up_d = __shfl_up_sync(0xffffffffu, d[j], 1), last_d = __shfl_sync(0xffffffffu, d[j], 31);
            const uint32_t up_i = __shfl_up_sync(0xffffffffu, id[j], 1), last_i = __shfl_sync(0xffffffffu, id[j], 31);
            if (pos < 32) {
                if ((int)lane > pos) { d[j] = up_d; id[j] = up_i; }
                else if ((int)lane == pos) { d[j] = nd; id[j] = nid; }
                nd = last_d; nid = last_i; shifting = true;
            }
            if (j * 32 + (int)lane >= cap) { d[j] = INFINITY; id[j] = NONE; }
        }
        refresh_worst();
    }

    // First unexpanded entry, not marked: the node the next pop() returns unless something closer is inserted first.
    __device__ __forceinline__ uint32_t peek() const {
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const unsigned m = __ballot_sync(0xffffffffu, id[j] != NONE && !(id[j] & EXPANDED));
            if (m) return __shfl_sync(0xffffffffu, id[j], __ffs(m) - 1);
        }
        return NONE;
    }

    // The first k live (not tombstoned) entries in ascending order -> caller keys + distances. Returns how many.
    __device__ __forceinline__ uint32_t emit_live(const GraphView& g, uint32_t kk, uint64_t* ko, float* dd) {
        const unsigned lane = lane_id();
        uint32_t nout = 0;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const uint32_t s = id[j] & ID_MASK;
            const bool live = id[j] != NONE && !is_deleted(g, s);
            const unsigned m = __ballot_sync(0xffffffffu, live);
            const uint32_t pos = nout + __popc(m & ((1u << lane) - 1u));
            if (live && pos < kk) { ko[pos] = g.slot_key[s]; dd[pos] = d[j]; }
            nout += __popc(m);
        }
        return nout > kk ? kk : nout;
    }

    // Every entry in ascending order -> ci / cd (efp entries, the tail padded with NONE / +inf); the list itself is kept.
    __device__ __forceinline__ uint32_t store_sorted(uint32_t* ci, float* cd, int efp) {
        const unsigned lane = lane_id();
        uint32_t n = 0;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const int i = j * 32 + (int)lane;
            if (i < efp) { ci[i] = id[j] == NONE ? NONE : (id[j] & ID_MASK); cd[i] = d[j]; }
            n += __popc(__ballot_sync(0xffffffffu, id[j] != NONE));
        }
        for (int i = NL * 32 + (int)lane; i < efp; i += 32) { ci[i] = NONE; cd[i] = INFINITY; }
        return n;
    }

    __device__ __forceinline__ void reopen_all(WarpSmem& S) {
#pragma unroll
        for (int j = 0; j < NL; ++j) if (id[j] != NONE) { id[j] &= ~EXPANDED; visited_ts(S.hs, S.hmask, S.tag16, id[j]); }
    }

    // First unexpanded entry; marks it expanded. NONE when the beam is finished.
    __device__ __forceinline__ uint32_t pop(float& d_out) {
        const unsigned lane = lane_id();
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const unsigned m = __ballot_sync(0xffffffffu, id[j] != NONE && !(id[j] & EXPANDED));
            if (m) {
                const int src = __ffs(m) - 1;
                const uint32_t pick = __shfl_sync(0xffffffffu, id[j], src);
                d_out = __shfl_sync(0xffffffffu, d[j], src);
                if ((int)lane == src) id[j] |= EXPANDED;
                return pick;
            }
        }
        return NONE;
    }
};

template <int NL>
struct UList {
    uint32_t k[NL];    // distance keys (KEY_NONE = empty)
    uint32_t id[NL];   // slot | EXPANDED | TAKEN, NONE = empty
    int cap, count;
    float worst;       // largest kept distance once the list is full, +inf before; warp-uniform
    int wlane, wrow;   // where that entry sits

    __device__ __forceinline__ void init(int cap_) {
        cap = cap_; count = 0; worst = INFINITY; wlane = 0; wrow = 0;
#pragma unroll
        for (int j = 0; j < NL; ++j) { k[j] = KEY_NONE; id[j] = NONE; }
    }

    __device__ __forceinline__ void set_at(int row, int lane, uint32_t kn, uint32_t nid) {
        const bool me = (int)lane_id() == lane;
#pragma unroll
        for (int j = 0; j < NL; ++j) if (j == row && me) { k[j] = kn; id[j] = nid; }
    }

    // The listed entry with the largest (distance, id): the one the next admitted candidate replaces.
    __device__ __forceinline__ void find_worst() {
        uint32_t mk = 0u, mi = 0u; int mj = 0;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const bool on = id[j] != NONE;
            const uint32_t kk = on ? k[j] : 0u, ii = on ? (id[j] & ID_MASK) : 0u;
            if (kk > mk || (kk == mk && ii >= mi)) { mk = kk; mi = ii; mj = j; }
        }
        const uint32_t w = __reduce_max_sync(0xffffffffu, mk);
        unsigned m = __ballot_sync(0xffffffffu, mk == w);
        if (m & (m - 1u)) {   // equal distances: the larger id is the worse one
            const uint32_t wi = __reduce_max_sync(0xffffffffu, mk == w ? mi : 0u);
            m = __ballot_sync(0xffffffffu, mk == w && mi == wi);
        }
        wlane = __ffs(m) - 1;
        wrow = __shfl_sync(0xffffffffu, mj, wlane);
        worst = fkey_inv(w);
    }

    // nd / nid are warp-uniform and nd < worst. Drops ids that are already listed.
    __device__ __forceinline__ void insert(float nd, uint32_t nid) {
        bool dup = false;
#pragma unroll
        for (int j = 0; j < NL; ++j) dup |= ((id[j] & ID_MASK) == nid) && id[j] != NONE;
        if (__any_sync(0xffffffffu, dup)) return;
        const uint32_t kn = fkey(nd);
        if (count < cap) {
            set_at(count >> 5, count & 31, kn, nid);
            if (++count == cap) find_worst();
        } else {
            set_at(wrow, wlane, kn, nid);
            find_worst();
        }
    }

    // The unexpanded entry with the smallest (distance, id); NONE when every entry is expanded. `mark`: flag it expanded.
    __device__ __forceinline__ uint32_t nearest_open(float& d_out, bool mark) {
        uint32_t bk = KEY_NONE, bi = NONE; int bj = 0;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const bool open = id[j] != NONE && !(id[j] & EXPANDED);
            const uint32_t kk = open ? k[j] : KEY_NONE, ii = id[j] & ID_MASK;
            if (kk < bk || (kk == bk && open && ii < bi)) { bk = kk; bi = ii; bj = j; }
        }
        const uint32_t w = __reduce_min_sync(0xffffffffu, bk);
        if (w == KEY_NONE) return NONE;
        unsigned m = __ballot_sync(0xffffffffu, bk == w);
        if (m & (m - 1u)) {
            const uint32_t wi = __reduce_min_sync(0xffffffffu, bk == w ? bi : NONE);
            m = __ballot_sync(0xffffffffu, bk == w && bi == wi);
        }
        const int src = __ffs(m) - 1;
        if (mark && (int)lane_id() == src) {
#pragma unroll
            for (int j = 0; j < NL; ++j) if (j == bj) id[j] |= EXPANDED;
        }
        d_out = fkey_inv(w);
        return __shfl_sync(0xffffffffu, bi, src);
    }
    __device__ __forceinline__ void seed(float nd, uint32_t nid) { insert(nd, nid); }
    __device__ __forceinline__ void reopen_all(WarpSmem& S) {
#pragma unroll
        for (int j = 0; j < NL; ++j) if (id[j] != NONE) { id[j] &= ~EXPANDED; visited_ts(S.hs, S.hmask, S.tag16, id[j] & ID_MASK); }
    }
    __device__ __forceinline__ uint32_t pop(float& d_out) { return nearest_open(d_out, true); }
    __device__ __forceinline__ uint32_t peek() { float d; return nearest_open(d, false); }

    // Ascending read-out: the not yet TAKEN entry with the smallest (distance, id), flagged TAKEN; NONE when none is left.
    __device__ __forceinline__ uint32_t take_min(float& d_out) {
        uint32_t bk = KEY_NONE, bi = NONE; int bj = 0;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const bool on = id[j] != NONE && !(id[j] & TAKEN);
            const uint32_t kk = on ? k[j] : KEY_NONE, ii = id[j] & ID_MASK;
            if (on && (kk < bk || (kk == bk && ii < bi))) { bk = kk; bi = ii; bj = j; }
        }
        const uint32_t w = __reduce_min_sync(0xffffffffu, bk);
        const unsigned any = __ballot_sync(0xffffffffu, bi != NONE);
        if (any == 0u) return NONE;
        unsigned m = __ballot_sync(0xffffffffu, bi != NONE && bk == w);
        if (m & (m - 1u)) {
            const uint32_t wi = __reduce_min_sync(0xffffffffu, (bi != NONE && bk == w) ? bi : NONE);
            m = __ballot_sync(0xffffffffu, bi != NONE && bk == w && bi == wi);
        }
        const int src = __ffs(m) - 1;
        if ((int)lane_id() == src) {
#pragma unroll
            for (int j = 0; j < NL; ++j) if (j == bj) id[j] |= TAKEN;
        }
        d_out = fkey_inv(w);
        return __shfl_sync(0xffffffffu, bi, src);
    }
    __device__ __forceinline__ void untake() {
#pragma unroll
        for (int j = 0; j < NL; ++j) if (id[j] != NONE) id[j] &= ~TAKEN;
    }

    // The first k live (not tombstoned) entries in ascending order -> caller keys + distances. Returns how many.
    __device__ __forceinline__ uint32_t emit_live(const GraphView& g, uint32_t kk, uint64_t* ko, float* dd) {
        uint32_t nout = 0;
        while (nout < kk) {
            float d; const uint32_t s = take_min(d);
            if (s == NONE) break;
            if (is_deleted(g, s)) continue;
            if (lane_id() == 0) { ko[nout] = g.slot_key[s]; dd[nout] = d; }
            ++nout;
        }
        return nout;
    }

    // Every entry in ascending order -> ci / cd (efp entries, the tail padded with NONE / +inf); the list itself is kept.
    __device__ __forceinline__ uint32_t store_sorted(uint32_t* ci, float* cd, int efp) {
        uint32_t n = 0;
        for (;;) {
            float d; const uint32_t s = take_min(d);
            if (s == NONE) break;
            if (lane_id() == 0) { ci[n] = s; cd[n] = d; }
            ++n;
        }
        for (int i = (int)n + (int)lane_id(); i < efp; i += 32) { ci[i] = NONE; cd[i] = INFINITY; }
        untake();
        return n;
    }
};

template <int NL> struct RListPick { typedef UList<NL> type; };
template <> struct RListPick<1> { typedef SortedList<1> type; };
template <> struct RListPick<2> { typedef SortedList<2> type; };
template <int NL> using RList = typename RListPick<NL>::type;

template <class LIST>
__device__ __forceinline__ void rbeam_level(const GraphView& g, int level, WarpSmem& S, LIST& L, float qn,
                                            uint32_t& evals, uint32_t& exps) {
    const unsigned lane = lane_id();
    for (;;) {
        float dc; const uint32_t c = L.pop(dc);
        if (c == NONE) break;
        ++exps;
        int width; const uint32_t* row = level_row(g, level, c, width);
        for (int off = 0; off < width; off += 32) {
            uint32_t nb = (off + (int)lane < width) ? row[off + lane] : NONE;
            if (nb != NONE && nb >= g.n_slots) nb = NONE;
            if (nb != NONE && visited_ts(S.hs, S.hmask, S.tag16, nb)) nb = NONE;
            __syncwarp();
            const float d = warp_dist32(g, S.q4, qn, nb, S.scr_id, S.scr_d, evals);
            unsigned acc = __ballot_sync(0xffffffffu, nb != NONE && d < L.worst);
            while (acc) {
                const int src = __ffs(acc) - 1; acc &= acc - 1;
                const float nd = __shfl_sync(0xffffffffu, d, src);
                const uint32_t nid = __shfl_sync(0xffffffffu, nb, src);
                if (nd < L.worst) L.insert(nd, nid);
            }
        }
    }
}

template <class LIST>
__device__ __forceinline__ void rbeam_seed(WarpSmem& S, LIST& L, int cap, uint32_t id, float d) {
    L.init(cap);
    visited_clear(S);
    if (lane_id() == 0) visited_ts(S.hs, S.hmask, S.tag16, id);
    __syncwarp();
    L.seed(d, id);
}

// Re-open every listed entry as an entry point of the next level down.
template <class LIST>
__device__ __forceinline__ void rbeam_reseed(WarpSmem& S, LIST& L) {
    visited_clear(S);
    L.reopen_all(S);
    __syncwarp();
}

}  // namespace vsann
