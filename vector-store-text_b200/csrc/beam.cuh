// beam.cuh — warp-per-query HNSW traversal (K3) shared by search and insertion.
//
// Restates, for one warp, the beam of usearch::Index::search / add that the reference calls at
// src/index/usearch.rs:276 and :221 (algorithm: SURVEY.md Appendix B; CPU restatement:
// oracle/oracle_vsann.cpp search_layer). The sorted list of the best `ef` candidates doubles as
// the candidate queue: the next node to expand is the first unexpanded list entry, and the beam
// ends when every entry of the list is expanded — the same stop rule as "closest candidate is
// farther than the worst of the top-ef".
#pragma once
#include "common.cuh"

namespace vsann {

struct WarpSmem {
    float4* q4;          // query, `chunks` float4
    SList L;             // ping-pong sorted lists
    uint32_t* hs;        // visited hash: 32-bit ids, or 16-bit tags two per word (tag16)
    uint32_t hmask;      // slots - 1
    uint32_t hwords;     // words behind hs
    uint32_t tag16;
    uint32_t* scr_id;    // 32
    float* scr_d;        // 32
};

// bytes of shared memory one warp needs
__host__ __device__ inline size_t warp_smem_bytes(uint32_t dpad, uint32_t capp, uint32_t hsize) {
    return (size_t)dpad * 4 + (size_t)capp * 16 + 256 + (size_t)hsize * 4;
}

__device__ __forceinline__ WarpSmem carve_warp_smem(unsigned char* base, uint32_t dpad, uint32_t cap, uint32_t capp, uint32_t hsize) {
    WarpSmem S;
    S.q4 = reinterpret_cast<float4*>(base); base += (size_t)dpad * 4;
    S.L.ld = reinterpret_cast<float*>(base); base += (size_t)capp * 4;
    S.L.ld2 = reinterpret_cast<float*>(base); base += (size_t)capp * 4;
    S.L.li = reinterpret_cast<uint32_t*>(base); base += (size_t)capp * 4;
    S.L.li2 = reinterpret_cast<uint32_t*>(base); base += (size_t)capp * 4;
    S.scr_id = reinterpret_cast<uint32_t*>(base); base += 128;
    S.scr_d = reinterpret_cast<float*>(base); base += 128;
    S.hs = reinterpret_cast<uint32_t*>(base);
    S.hmask = hsize - 1; S.hwords = hsize; S.tag16 = 0;
    S.L.cap = cap; S.L.capp = capp;
    return S;
}

__device__ __forceinline__ void visited_clear(WarpSmem& S) {
    const uint32_t empty = S.tag16 ? 0u : NONE;
    for (uint32_t i = lane_id(); i < S.hwords; i += 32) S.hs[i] = empty;
    __syncwarp();
}
__device__ __forceinline__ bool visited_ts(uint32_t* hs, uint32_t hmask, uint32_t tag16, uint32_t id) {
    return tag16 ? visited_test_set16(hs, hmask, id) : visited_test_set(hs, hmask, id);
}

// Greedy 1-best walk on `level` (upper levels). cur/dcur in, out.
__device__ __forceinline__ void greedy_level(const GraphView& g, int level, WarpSmem& S, float qn,
                                             uint32_t& cur, float& dcur, uint32_t& evals, uint32_t& exps) {
    const unsigned lane = lane_id();
    for (;;) {
        int width; const uint32_t* row = level_row(g, level, cur, width);
        ++exps;
        float bd = INFINITY; uint32_t bi = NONE;
        for (int off = 0; off < width; off += 32) {
            uint32_t nb = (off + (int)lane < width) ? row[off + lane] : NONE;
            if (nb != NONE && nb >= g.n_slots) nb = NONE;
            const float d = warp_dist32(g, S.q4, qn, nb, S.scr_id, S.scr_d, evals);
            if (nb != NONE && pair_less(d, nb, bd, bi)) { bd = d; bi = nb; }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, bd, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (pair_less(od, oi, bd, bi)) { bd = od; bi = oi; }
        }
        if (bi != NONE && bd < dcur) { cur = bi; dcur = bd; } else break;
    }
}

// ef-wide beam on `level`. The list must already hold the (unexpanded) entry points and the
// visited set must contain them.
__device__ __forceinline__ void beam_level(const GraphView& g, int level, WarpSmem& S, float qn,
                                           uint32_t& evals, uint32_t& exps) {
    const unsigned lane = lane_id();
    for (;;) {
        float dc; const uint32_t c = slist_pop(S.L, dc);
        if (c == NONE) break;
        ++exps;
        int width; const uint32_t* row = level_row(g, level, c, width);
        for (int off = 0; off < width; off += 32) {
            uint32_t nb = (off + (int)lane < width) ? row[off + lane] : NONE;
            if (nb != NONE && nb >= g.n_slots) nb = NONE;
            if (nb != NONE && visited_ts(S.hs, S.hmask, S.tag16, nb)) nb = NONE;
            __syncwarp();
            float d = warp_dist32(g, S.q4, qn, nb, S.scr_id, S.scr_d, evals);
            const float worst = slist_worst(S.L);
            if (!(d < worst)) { nb = NONE; d = INFINITY; }
            slist_merge(S.L, d, nb, S.scr_id, S.scr_d);
        }
    }
}

// Put one entry point into a cleared list / visited set.
__device__ __forceinline__ void beam_seed(WarpSmem& S, uint32_t id, float d) {
    slist_init(S.L);
    visited_clear(S);
    if (lane_id() == 0) { S.L.ld[0] = d; S.L.li[0] = id; visited_ts(S.hs, S.hmask, S.tag16, id); }
    __syncwarp();
}

// Re-open the best entries of the current list as entry points for the next level down:
// clears the expanded flags and the visited set, re-marks the listed ids.
__device__ __forceinline__ void beam_reseed_from_list(WarpSmem& S) {
    visited_clear(S);
    for (int i = lane_id(); i < S.L.cap; i += 32) {
        uint32_t id = S.L.li[i];
        if (id != NONE) { id &= ~EXPANDED; S.L.li[i] = id; visited_ts(S.hs, S.hmask, S.tag16, id); }
    }
    __syncwarp();
}

__device__ __forceinline__ float warp_entry_dist(const GraphView& g, WarpSmem& S, float qn, uint32_t entry, uint32_t& evals) {
    const float d = warp_dist32(g, S.q4, qn, lane_id() == 0 ? entry : NONE, S.scr_id, S.scr_d, evals);
    return __shfl_sync(0xffffffffu, d, 0);
}

}  // namespace vsann
