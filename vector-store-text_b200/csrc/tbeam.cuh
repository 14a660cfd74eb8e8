// tbeam.cuh — bulk-copy (TMA) staged gathers for the warp beam.
//
// Same traversal as rbeam.cuh (reference: usearch::Index::search / add, src/index/usearch.rs:276,221),
// but the neighbour vectors of one expansion are fetched by ONE round of `cp.async.bulk`
// (global -> shared, one row per lane, completion on a per-warp mbarrier) instead of four
// dependent rounds of register loads: a warp keeps up to `rows` x row_bytes in flight without
// holding them in registers, so a beam step costs two dependent memory round trips (link row,
// then all vectors) and the link row of every accepted candidate is prefetched into L2 as soon
// as it is accepted. Distances: 8-lane teams, chunk c = tl + 8 i per lane on the packed-f32 pipe (even / odd
// partial sums, acc_chunk in common.cuh), xor-4/2/1 reduction: the same bits as warp_dist32.
#pragma once
#include <cuda.h>
#include "rbeam.cuh"

namespace vsann {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
// global -> shared bulk copy (bytes % 16 == 0, both addresses 16 B aligned), completes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// 16-byte asynchronous copy global -> shared (LDGSTS), bypassing L1; completion by commit/wait group
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// four rows of a 2-D tensor (arbitrary row indices, box = one row) -> four consecutive rows in shared memory
__device__ __forceinline__ void tma_gather4(uint32_t dst, const CUtensorMap* map, uint32_t bar, int r0, int r1, int r2, int r3) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}
// one instruction brings `bytes` (multiple of 16) into L2; nothing is waited for
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// read-only probe of the visited hash (same probe sequence as visited_test_set)
__device__ __forceinline__ bool visited_probe(const uint32_t* hs, uint32_t hmask, uint32_t id) {
    const uint32_t h = (id * 0x9E3779B1u) >> 7;
#pragma unroll 1
    for (int probe = 0; probe < 8; ++probe) {
        const uint32_t v = hs[(h + probe) & hmask];
        if (v == id) return true;
        if (v == NONE) return false;
    }
    return false;
}
__device__ __forceinline__ void prefetch_l2(const void* p) {
#ifndef VSANN_NO_PREFETCH
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}

// rows longer than this use cp.async.bulk, shorter ones LDGSTS. Measured on B200 (1M x 128, ef=24):
// bulk 11.4M QPS vs LDGSTS 9.6M QPS, so every row goes through the bulk path.
constexpr uint32_t BULK_MIN_ROW_BYTES = 0;
#ifndef VSANN_PF_NEXT
#define VSANN_PF_NEXT 0
#endif
// L2 prefetch (UBLKPF) of the likely next node's unvisited neighbours while the current gather is in flight. Measured on
// B200 (1M x 128, ef=24): 13.7M QPS with it vs 15.2M without, and slower even for a single wave of queries (0.408 vs
// 0.377 ms at 2500 queries): the probe + prefetch issue costs more than the L2 hits return. Off.
constexpr bool PF_NEXT = VSANN_PF_NEXT != 0;

struct WarpStage {
    const unsigned char* buf;   // rows x row_bytes staging area (generic pointer)
    uint32_t buf_s;             // its shared-space address
    uint32_t bar;               // shared-space address of the warp's mbarriers (slot s: bar + 8 s)
    uint32_t parity;            // bit s: parity of slot s's next completion
    uint32_t rows;              // staging capacity in rows (a multiple of 4)
    uint32_t row_bytes;         // dpad * 4
    const CUtensorMap* tmap;    // vecs as a [slots][dpad] tensor with a one-row box (null: rows longer than a TMA box)
    bool pf_over;               // L2-prefetch the rows of an expansion that do not fit the staging area
};

// per-warp layout: query | scr_id[32] | scr_d[32] | 8 mbarriers (one per staging slot of four rows) | visited hash | staging rows (128 B aligned)
__host__ __device__ inline size_t tsmem_bytes(uint32_t dpad, uint32_t hsize, uint32_t rows) {
    size_t b = (size_t)dpad * 4 + 256 + 64 + (size_t)hash_slots(hsize) * (hash_uses_tags(hsize) ? 2 : 4);
    b = (b + 127) & ~(size_t)127;
    return b + (size_t)rows * dpad * 4;
}

__device__ __forceinline__ void carve_tsmem(unsigned char* base, uint32_t dpad, uint32_t hsize, uint32_t rows, WarpSmem& S, WarpStage& T) {
    unsigned char* p = base;
    S.q4 = reinterpret_cast<float4*>(p); p += (size_t)dpad * 4;
    S.scr_id = reinterpret_cast<uint32_t*>(p); p += 128;
    S.scr_d = reinterpret_cast<float*>(p); p += 128;
    T.bar = smem_u32(p); p += 64;
    S.tag16 = hash_uses_tags(hsize) ? 1u : 0u; S.hwords = S.tag16 ? hash_slots(hsize) / 2 : hash_slots(hsize);
    S.hs = reinterpret_cast<uint32_t*>(p); p += (size_t)S.hwords * 4;
    S.hmask = hash_slots(hsize) - 1;
    p = base + ((((size_t)(p - base)) + 127) & ~(size_t)127);
    T.buf = p; T.buf_s = smem_u32(p);
    T.rows = rows; T.row_bytes = dpad * 4; T.parity = 0; T.tmap = nullptr; T.pf_over = false;
    S.L.ld = S.L.ld2 = nullptr; S.L.li = S.L.li2 = nullptr; S.L.cap = S.L.capp = 0;
    if (lane_id() < 8) mbar_init(T.bar + 8u * lane_id(), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
}

// Query chunks a lane owns in the team layout: c = tl + 8 i. CPL == 0: read them from shared memory.
template <int CPL>
struct QRegs {
    float4 v[CPL > 0 ? CPL : 1];
    __device__ __forceinline__ void load(const float4* q4) {
        const unsigned tl = lane_id() & 7u;
#pragma unroll
        for (int i = 0; i < CPL; ++i) v[i] = q4[tl + 8 * i];
    }
};

// One group = four rows into staging slot `slot` (4 x row_bytes), completion on that slot's mbarrier. Called by ONE lane.
// With a tensor map it is a single TMA gather4 (four scattered rows per instruction), else four bulk copies.
__device__ __forceinline__ void stage_issue_group(const GraphView& g, const WarpSmem& S, const WarpStage& T, int grp, int slot) {
    const uint4 r = *reinterpret_cast<const uint4*>(S.scr_id + 4 * grp);
    const uint32_t bar = T.bar + 8u * (uint32_t)slot, dst = T.buf_s + (uint32_t)slot * 4u * T.row_bytes;
    mbar_arrive_expect_tx(bar, 4u * T.row_bytes);
    if (T.tmap) {
        tma_gather4(dst, T.tmap, bar, (int)r.x, (int)r.y, (int)r.z, (int)r.w);
    } else {
        bulk_g2s(dst, g.vecs + (size_t)r.x * g.dpad, T.row_bytes, bar);
        bulk_g2s(dst + T.row_bytes, g.vecs + (size_t)r.y * g.dpad, T.row_bytes, bar);
        bulk_g2s(dst + 2u * T.row_bytes, g.vecs + (size_t)r.z * g.dpad, T.row_bytes, bar);
        bulk_g2s(dst + 3u * T.row_bytes, g.vecs + (size_t)r.w * g.dpad, T.row_bytes, bar);
    }
}

// Distances from the warp's query to up to 32 rows, one id per lane (NONE = skip); +inf for skipped lanes.
// The rows stream through the staging area in groups of four (one per 8-lane team), rows/4 groups in flight:
// while a group's distances are computed the next groups are still landing, and a freed slot is refilled at once,
// so an expansion with more rows than the staging area holds pays the memory latency once, not once per refill.
template <int METRIC, int CPL>
__device__ __forceinline__ float stage_dist32_m(const GraphView& g, WarpSmem& S, WarpStage& T, const QRegs<CPL>& Q, float qn,
                                                uint32_t id, uint32_t& n_evals, uint32_t pf = NONE) {
    const unsigned lane = lane_id(), team = lane >> 3, tl = lane & 7u;
    const unsigned valid = __ballot_sync(0xffffffffu, id != NONE);
    if (valid == 0u) {
        if (pf != NONE && !S.tag16 && pf < g.n_slots && !visited_probe(S.hs, S.hmask, pf)) bulk_prefetch_l2(g.vecs + (size_t)pf * g.dpad, T.row_bytes);
        return INFINITY;
    }
    const int nvalid = __popc(valid);
    const int pos = __popc(valid & ((1u << lane) - 1u));
    const uint32_t first = __shfl_sync(0xffffffffu, id, __ffs(valid) - 1);
    if (id != NONE) S.scr_id[pos] = id;
    if ((int)lane >= nvalid) S.scr_id[lane] = first;   // pad the last group with a row that is fetched anyway
    __syncwarp();
    n_evals += nvalid;
    const int chunks = g.chunks;
    const int ngroups = (nvalid + 3) >> 2, ns = (int)(T.rows >> 2);
    if ((int)lane < min(ns, ngroups)) stage_issue_group(g, S, T, (int)lane, (int)lane);
    // rows beyond the staging capacity wait for a free slot: have them in L2 by then (they are needed for certain).
    // Measured on B200: search 15.15M -> 15.73M QPS; the build search (ef_add wide, mostly L2-resident) loses 2 %, so it is per kernel.
    if (T.pf_over && (int)lane >= (int)T.rows && (int)lane < nvalid) bulk_prefetch_l2(g.vecs + (size_t)S.scr_id[lane] * g.dpad, T.row_bytes);
    // while this expansion's rows are in flight: bring the not yet visited neighbours of the likely next node into L2
    if (pf != NONE && !S.tag16 && pf < g.n_slots && !visited_probe(S.hs, S.hmask, pf)) bulk_prefetch_l2(g.vecs + (size_t)pf * g.dpad, T.row_bytes);
    if (METRIC == COS) { if (id != NONE) S.scr_d[pos] = g.sqn[id]; __syncwarp(); }
    int slot = 0;
    for (int grp = 0; grp < ngroups; ++grp) {
        mbar_wait(T.bar + 8u * (uint32_t)slot, (T.parity >> slot) & 1u); T.parity ^= 1u << slot;
        const int j = 4 * grp + (int)team;
        const float4* x4 = reinterpret_cast<const float4*>(T.buf + ((size_t)slot * 4u + team) * T.row_bytes);
        float2 acc2 = make_float2(0.f, 0.f);
        if (CPL > 0) {
            float4 x[CPL > 0 ? CPL : 1];
#pragma unroll
            for (int i = 0; i < CPL; ++i) x[i] = x4[tl + 8 * i];
#pragma unroll
            for (int i = 0; i < CPL; ++i) acc_chunk<METRIC>(Q.v[i], x[i], acc2);
        } else {
#pragma unroll 4
            for (int c = tl; c < chunks; c += 8) acc_chunk<METRIC>(S.q4[c], x4[c], acc2);
        }
        float acc = acc2.x + acc2.y;
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (tl == 0 && j < nvalid) S.scr_d[j] = finish_metric(METRIC, acc, qn, METRIC == COS ? S.scr_d[j] : 0.f);
        __syncwarp();   // the slot's rows are consumed, distances visible
        if (grp + ns < ngroups && lane == 0) stage_issue_group(g, S, T, grp + ns, slot);
        if (++slot == ns) slot = 0;
    }
    const float d = id != NONE ? S.scr_d[pos] : INFINITY;
    __syncwarp();
    return d;
}

template <int CPL>
__device__ __forceinline__ float stage_dist32(const GraphView& g, WarpSmem& S, WarpStage& T, const QRegs<CPL>& Q, float qn,
                                              uint32_t id, uint32_t& n_evals, uint32_t pf = NONE) {
    switch (g.metric) {
        case L2SQ: return stage_dist32_m<L2SQ, CPL>(g, S, T, Q, qn, id, n_evals, pf);
        case IP: return stage_dist32_m<IP, CPL>(g, S, T, Q, qn, id, n_evals, pf);
        default: return stage_dist32_m<COS, CPL>(g, S, T, Q, qn, id, n_evals, pf);
    }
}

template <int CPL>
__device__ __forceinline__ float stage_entry_dist(const GraphView& g, WarpSmem& S, WarpStage& T, const QRegs<CPL>& Q, float qn,
                                                  uint32_t entry, uint32_t& evals) {
    const float d = stage_dist32<CPL>(g, S, T, Q, qn, lane_id() == 0 ? entry : NONE, evals);
    return __shfl_sync(0xffffffffu, d, 0);
}

// Greedy 1-best walk on an upper level (cf. greedy_level in beam.cuh).
template <int CPL>
__device__ __forceinline__ void tgreedy_level(const GraphView& g, int level, WarpSmem& S, WarpStage& T, const QRegs<CPL>& Q, float qn,
                                              uint32_t& cur, float& dcur, uint32_t& evals, uint32_t& exps) {
    const unsigned lane = lane_id();
    for (;;) {
        int width; const uint32_t* row = level_row(g, level, cur, width);
        ++exps;
        float bd = INFINITY; uint32_t bi = NONE;
        for (int off = 0; off < width; off += 32) {
            uint32_t nb = (off + (int)lane < width) ? row[off + lane] : NONE;
            if (nb != NONE && nb >= g.n_slots) nb = NONE;
            const float d = stage_dist32<CPL>(g, S, T, Q, qn, nb, evals);
            if (nb != NONE && pair_less(d, nb, bd, bi)) { bd = d; bi = nb; }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, bd, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (pair_less(od, oi, bd, bi)) { bd = od; bi = oi; }
        }
        if (bi != NONE && bd < dcur) {
            cur = bi; dcur = bd;
            int w2; prefetch_l2(level_row(g, level, cur, w2));
        } else break;
    }
}

// ef-wide beam on `level` with the register list (cf. rbeam_level).
template <int NL, int CPL>
__device__ __forceinline__ void tbeam_level(const GraphView& g, int level, WarpSmem& S, WarpStage& T, const QRegs<CPL>& Q,
                                            RList<NL>& L, float qn, uint32_t& evals, uint32_t& exps) {
    const unsigned lane = lane_id();
    // The link row of the most likely NEXT node is loaded while the current one is expanded (a row is at most
    // 32 links: one per lane, one register): when the guess holds, a beam step has one dependent memory round
    // trip (the vector gather) instead of two.
    uint32_t spec_id = NONE, spec_nb = NONE;
    for (;;) {
        float dc; const uint32_t c = L.pop(dc);
        if (c == NONE) break;
        ++exps;
        int width; const uint32_t* row = level_row(g, level, c, width);
        uint32_t nb_row = spec_nb;
        if (c != spec_id) nb_row = ((int)lane < width) ? row[lane] : NONE;
        spec_id = L.peek();
        if (spec_id != NONE) { int w2; const uint32_t* r2 = level_row(g, level, spec_id, w2); spec_nb = ((int)lane < w2) ? r2[lane] : NONE; }
        {
            uint32_t nb = nb_row;
            if (nb != NONE && nb >= g.n_slots) nb = NONE;
            if (nb != NONE && visited_ts(S.hs, S.hmask, S.tag16, nb)) nb = NONE;
            __syncwarp();
            const float d = stage_dist32<CPL>(g, S, T, Q, qn, nb, evals, PF_NEXT && spec_id != NONE ? spec_nb : NONE);
            const bool good = nb != NONE && d < L.worst;
            if (good) { int w2; prefetch_l2(level_row(g, level, nb, w2)); }   // it may be expanded later: have its links in L2
            unsigned acc = __ballot_sync(0xffffffffu, good);
            while (acc) {
                const int src = __ffs(acc) - 1; acc &= acc - 1;
                const float nd = __shfl_sync(0xffffffffu, d, src);
                const uint32_t nid = __shfl_sync(0xffffffffu, nb, src);
                if (nd < L.worst) L.insert(nd, nid);
            }
        }
    }
}

}  // namespace vsann
