// rbeam.cuh — register-resident candidate list for the warp beam (ef <= 32 * NL).
//
// The sorted list of the best `cap` candidates lives in registers, striped across the warp:
// entry i sits in row i / 32 of lane i % 32. Popping the next node is one ballot per row;
// inserting a candidate is one ballot + one shuffle-up per row. Same semantics as SList
// (beam.cuh): equivalent to the oracle's two-heap beam (oracle_vsann.cpp search_layer).
#pragma once
#include "beam.cuh"

namespace vsann {

template <int NL>
struct RList {
    float d[NL];
    uint32_t id[NL];
    int cap;
    float worst;   // distance of entry cap-1 (+inf while the list is not full), warp-uniform

    __device__ __forceinline__ void init(int cap_) {
        cap = cap_; worst = INFINITY;
#pragma unroll
        for (int j = 0; j < NL; ++j) { d[j] = INFINITY; id[j] = NONE; }
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
__device__ __forceinline__ void rbeam_level(const GraphView& g, int level, WarpSmem& S, RList<NL>& L, float qn,
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
            if (nb != NONE && visited_test_set(S.hs, S.hmask, nb)) nb = NONE;
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

template <int NL>
__device__ __forceinline__ void rbeam_seed(WarpSmem& S, RList<NL>& L, int cap, uint32_t id, float d) {
    L.init(cap);
    visited_clear(S);
    if (lane_id() == 0) { L.d[0] = d; L.id[0] = id; visited_test_set(S.hs, S.hmask, id); }
    __syncwarp();
    L.refresh_worst();
}

// Re-open every listed entry as an entry point of the next level down.
template <int NL>
__device__ __forceinline__ void rbeam_reseed(WarpSmem& S, RList<NL>& L) {
    visited_clear(S);
#pragma unroll
    for (int j = 0; j < NL; ++j) if (L.id[j] != NONE) { L.id[j] &= ~EXPANDED; visited_test_set(S.hs, S.hmask, L.id[j]); }
    __syncwarp();
}

}  // namespace vsann
