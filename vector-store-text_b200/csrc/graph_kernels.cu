// graph_kernels.cu — K3 (ANN search), K4 (batched insertion: search + heuristic select) and
// K5 (reverse-link merge / re-prune) for the fixed-degree neighbour tables.
//
// Reference behaviour restated: usearch::Index::search (src/index/usearch.rs:276) and
// usearch::Index::add (src/index/usearch.rs:221); algorithm per SURVEY.md Appendix B, CPU
// restatement in oracle/oracle_vsann.cpp (hnsw_search / hnsw_insert / select_heuristic).
#include <algorithm>
#include <atomic>
#include <cstring>
#include "tbeam.cuh"
#include "kernels.h"
#include "tmap.h"

#ifndef VSANN_T_MINB
#define VSANN_T_MINB 6   // resident blocks per SM the staged kernels are compiled for
#endif

namespace vsann {

__device__ __forceinline__ float warp_sqnorm(const float4* q4, int chunks) {
    float acc = 0.f;
    for (int c = lane_id(); c < chunks; c += 32) { const float4 v = q4[c]; acc = fmaf(v.x, v.x, acc); acc = fmaf(v.y, v.y, acc); acc = fmaf(v.z, v.z, acc); acc = fmaf(v.w, v.w, acc); }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    return acc;
}

// Loads a query given with `dims` floats (row stride `ld`) into the padded shared buffer.
__device__ __forceinline__ void warp_load_query(float4* q4, const float* q, uint32_t dims, uint32_t dpad) {
    float* qs = reinterpret_cast<float*>(q4);
    for (uint32_t i = lane_id(); i < dpad; i += 32) qs[i] = i < dims ? q[i] : 0.f;
    __syncwarp();
}

// ---------------------------------------------------------------- K3: ANN search ------------
__global__ void __launch_bounds__(128, 8) k_search(GraphView g, const float* __restrict__ queries, uint32_t nq, uint32_t k,
                                                uint32_t ef, uint32_t efp, uint32_t hsize,
                                                uint64_t* __restrict__ keys_out, float* __restrict__ dist_out,
                                                unsigned long long* counters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned wpb = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = warp_smem_bytes(g.dpad, efp, hsize);
    WarpSmem S = carve_warp_smem(smem + warp * per_warp, g.dpad, ef, efp, hsize);
    uint32_t evals = 0, exps = 0, done = 0;
    for (uint32_t q = blockIdx.x * wpb + warp; q < nq; q += gridDim.x * wpb) {
        warp_load_query(S.q4, queries + (size_t)q * g.dims, g.dims, g.dpad);
        uint64_t* ko = keys_out + (size_t)q * k; float* dd = dist_out + (size_t)q * k;
        uint32_t nout = 0;
        if (g.n_slots > 0) {
            const float qn = g.metric == COS ? warp_sqnorm(S.q4, g.chunks) : 0.f;
            uint32_t cur = 0; float dcur = warp_entry_dist(g, S, qn, 0u, evals);
            for (int l = g.top_level; l >= 1; --l) greedy_level(g, l, S, qn, cur, dcur, evals, exps);
            beam_seed(S, cur, dcur);
            beam_level(g, 0, S, qn, evals, exps);
            // emit the first k live entries
            for (int base = 0; base < (int)ef && nout < k; base += 32) {
                const int i = base + lane;
                uint32_t id = i < (int)ef ? S.L.li[i] : NONE;
                const float d = i < (int)ef ? S.L.ld[i] : INFINITY;
                bool live = false;
                if (id != NONE) { id &= ~EXPANDED; live = !is_deleted(g, id); }
                const unsigned m = __ballot_sync(0xffffffffu, live);
                const uint32_t pos = nout + __popc(m & ((1u << lane) - 1u));
                if (live && pos < k) { ko[pos] = g.slot_key[id]; dd[pos] = d; }
                nout += __popc(m);
            }
            if (nout > k) nout = k;
        }
        for (uint32_t i = nout + lane; i < k; i += 32) { ko[i] = 0xFFFFFFFFFFFFFFFFull; dd[i] = INFINITY; }
        ++done;
        __syncwarp();
    }
    if (counters && lane == 0) {
        atomicAdd(&counters[0], (unsigned long long)done);
        atomicAdd(&counters[1], (unsigned long long)evals);
        atomicAdd(&counters[2], (unsigned long long)exps);
    }
}

// Register-list variant (ef <= 32 * NL): the hot configuration of K3.
template <int NL>
__global__ void __launch_bounds__(128, 8) k_search_r(GraphView g, const float* __restrict__ queries, uint32_t nq, uint32_t k,
                                                     uint32_t ef, uint32_t hsize,
                                                     uint64_t* __restrict__ keys_out, float* __restrict__ dist_out,
                                                     unsigned long long* counters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned wpb = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = warp_smem_bytes(g.dpad, 0, hsize);
    WarpSmem S = carve_warp_smem(smem + warp * per_warp, g.dpad, 0, 0, hsize);
    uint32_t evals = 0, exps = 0, done = 0;
    for (uint32_t q = blockIdx.x * wpb + warp; q < nq; q += gridDim.x * wpb) {
        warp_load_query(S.q4, queries + (size_t)q * g.dims, g.dims, g.dpad);
        uint64_t* ko = keys_out + (size_t)q * k; float* dd = dist_out + (size_t)q * k;
        uint32_t nout = 0;
        if (g.n_slots > 0) {
            const float qn = g.metric == COS ? warp_sqnorm(S.q4, g.chunks) : 0.f;
            uint32_t cur = 0; float dcur = warp_entry_dist(g, S, qn, 0u, evals);
            for (int l = g.top_level; l >= 1; --l) greedy_level(g, l, S, qn, cur, dcur, evals, exps);
            RList<NL> L;
            rbeam_seed(S, L, (int)ef, cur, dcur);
            rbeam_level(g, 0, S, L, qn, evals, exps);
            nout = L.emit_live(g, k, ko, dd);
        }
        for (uint32_t i = nout + lane; i < k; i += 32) { ko[i] = 0xFFFFFFFFFFFFFFFFull; dd[i] = INFINITY; }
        ++done;
        __syncwarp();
    }
    if (counters && lane == 0) {
        atomicAdd(&counters[0], (unsigned long long)done);
        atomicAdd(&counters[1], (unsigned long long)evals);
        atomicAdd(&counters[2], (unsigned long long)exps);
    }
}

// Copies of a query's finished result list into up to 8 more buffers - other GPUs' memory mapped over NVLink / NVSwitch in the sharded
// search (vs_search_batch_fanout_device): the lists travel while the traversal of the other queries is still running, and the merge
// that follows reads local memory only. The warp re-reads what it has just written (ordered by the __syncwarp before the call).
__device__ __forceinline__ void fan_out(const OutFan& fan, uint32_t q, uint32_t k, const uint64_t* ko, const float* dd) {
    if (fan.n == 0) return;
    for (uint32_t i = lane_id(); i < k; i += 32) {
        const uint64_t key = ko[i]; const float d = dd[i];
        for (uint32_t r = 0; r < fan.n; ++r) { fan.keys[r][(size_t)q * k + i] = key; fan.dist[r][(size_t)q * k + i] = d; }
    }
}

// Bulk-copy staged variant (tbeam.cuh): the default K3. NL = list rows (ef <= 32 NL), CPL = query
// float4 chunks a lane keeps in registers (chunks == 8 CPL), 0 = any dims (query read from shared memory).
template <int NL, int CPL>
__global__ void __launch_bounds__(128, (NL <= 8 ? VSANN_T_MINB : (NL <= 16 ? 4 : 3))) k_search_t(const __grid_constant__ CUtensorMap vmap, int use_vmap, GraphView g, const float* __restrict__ queries, uint32_t nq, uint32_t k,
                                                     uint32_t ef, uint32_t hsize, uint32_t stage_rows,
                                                     uint64_t* __restrict__ keys_out, float* __restrict__ dist_out,
                                                     unsigned long long* counters, uint32_t* work, const OutFan fan) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = tsmem_bytes(g.dpad, hsize, stage_rows);
    WarpSmem S; WarpStage T;
    carve_tsmem(smem + warp * per_warp, g.dpad, hsize, stage_rows, S, T);
    if (use_vmap) T.tmap = &vmap;
    T.pf_over = true;
    uint32_t evals = 0, exps = 0, done = 0;
    for (;;) {   // persistent warps: every warp draws its next query from a device counter (no tail, no intra-block wait)
        uint32_t q = 0;
        if (lane == 0) q = atomicAdd(work, 1u);
        q = __shfl_sync(0xffffffffu, q, 0);
        if (q >= nq) break;
        warp_load_query(S.q4, queries + (size_t)q * g.dims, g.dims, g.dpad);
        QRegs<CPL> Q; Q.load(S.q4);
        uint64_t* ko = keys_out + (size_t)q * k; float* dd = dist_out + (size_t)q * k;
        uint32_t nout = 0;
        if (g.n_slots > 0) {
            const float qn = g.metric == COS ? warp_sqnorm(S.q4, g.chunks) : 0.f;
            uint32_t cur = 0; float dcur = stage_entry_dist<CPL>(g, S, T, Q, qn, 0u, evals);
            for (int l = g.top_level; l >= 1; --l) tgreedy_level<CPL>(g, l, S, T, Q, qn, cur, dcur, evals, exps);
            RList<NL> L;
            rbeam_seed(S, L, (int)ef, cur, dcur);
            tbeam_level<NL, CPL>(g, 0, S, T, Q, L, qn, evals, exps);
            nout = L.emit_live(g, k, ko, dd);
        }
        for (uint32_t i = nout + lane; i < k; i += 32) { ko[i] = 0xFFFFFFFFFFFFFFFFull; dd[i] = INFINITY; }
        ++done;
        __syncwarp();
        fan_out(fan, q, k, ko, dd);
    }
    if (counters && lane == 0) {
        atomicAdd(&counters[0], (unsigned long long)done);
        atomicAdd(&counters[1], (unsigned long long)evals);
        atomicAdd(&counters[2], (unsigned long long)exps);
    }
}

// ------------------------------------------------- K3b: one BLOCK per query (small batches) ----
// With a handful of queries a warp per query leaves the GPU idle and a query waits for one warp to stream its
// neighbour rows (5M x 1536, ef = 128: ~10 MB through one warp's 24 KB staging area = 1.1 ms per query). Here a block
// of W warps serves one query: warp 0 (the leader) owns the candidate list, the visited hash and the link rows exactly
// as in k_search_t; the up to 32 unvisited neighbours of an expansion are split among ALL warps, each warp streams
// its share through its own staging area (the same stage_dist32 code, so the same distance bits), and the leader
// continues with the distances. Two block barriers per expansion. Same lists as k_search_t, query by query.
struct BlkShared { int n; uint32_t id[32]; float d[32]; };

template <int CPL>
__device__ __forceinline__ float blk_dist32(const GraphView& g, WarpSmem& S, WarpStage& T, const QRegs<CPL>& Q, float qn, BlkShared* B,
                                            uint32_t nb /* leader lanes: ids or NONE */, bool go /* leader: false = finished */, bool& finished,
                                            uint32_t& evals) {
    const unsigned warp = threadIdx.x >> 5, lane = lane_id(), W = blockDim.x >> 5;
    int pos = 0;
    if (warp == 0) {
        const unsigned valid = __ballot_sync(0xffffffffu, go && nb != NONE);
        pos = __popc(valid & ((1u << lane) - 1u));
        if (go && nb != NONE) B->id[pos] = nb;
        if (lane == 0) B->n = go ? __popc(valid) : -1;
    }
    __syncthreads();
    const int n = B->n;
    finished = n < 0;
    if (finished) return INFINITY;
    const int R = 4 * ((n + 4 * (int)W - 1) / (4 * (int)W));            // ids per warp: whole gather4 groups
    const int mine = (int)warp * R + (int)lane;
    const uint32_t id = (R > 0 && (int)lane < R && mine < n) ? B->id[mine] : NONE;
    uint32_t ev = 0;
    const float d = stage_dist32<CPL>(g, S, T, Q, qn, id, ev);
    if (id != NONE) B->d[mine] = d;
    __syncthreads();
    if (warp == 0) evals += (uint32_t)n;
    return (warp == 0 && nb != NONE) ? B->d[pos] : INFINITY;
}

template <int NL, int CPL>
__global__ void __launch_bounds__(256) k_search_blk(const __grid_constant__ CUtensorMap vmap, int use_vmap, GraphView g, const float* __restrict__ queries,
                                                    uint32_t nq, uint32_t k, uint32_t ef, uint32_t hsize, uint32_t stage_rows,
                                                    uint64_t* __restrict__ keys_out, float* __restrict__ dist_out, unsigned long long* counters, const OutFan fan) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ BlkShared B;
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = tsmem_bytes(g.dpad, hsize, stage_rows);
    WarpSmem S; WarpStage T;
    carve_tsmem(smem + warp * per_warp, g.dpad, hsize, stage_rows, S, T);
    if (use_vmap) T.tmap = &vmap;
    T.pf_over = true;
    const uint32_t q = blockIdx.x;
    if (q >= nq) return;
    uint32_t evals = 0, exps = 0;
    warp_load_query(S.q4, queries + (size_t)q * g.dims, g.dims, g.dpad);      // every warp keeps its own copy of the query
    QRegs<CPL> Q; Q.load(S.q4);
    uint64_t* ko = keys_out + (size_t)q * k; float* dd = dist_out + (size_t)q * k;
    uint32_t nout = 0;
    if (g.n_slots > 0) {
        const float qn = g.metric == COS ? warp_sqnorm(S.q4, g.chunks) : 0.f;
        bool fin;
        // entry point, then the greedy descent (leader decides, everybody evaluates)
        uint32_t cur = 0;
        float dcur = blk_dist32<CPL>(g, S, T, Q, qn, &B, lane == 0 ? 0u : NONE, true, fin, evals);
        dcur = __shfl_sync(0xffffffffu, dcur, 0);
        for (int l = g.top_level; l >= 1; --l) {
            for (;;) {
                uint32_t nb = NONE;
                if (warp == 0) {
                    int width; const uint32_t* row = level_row(g, l, cur, width);
                    ++exps;
                    nb = ((int)lane < width) ? row[lane] : NONE;
                    if (nb != NONE && nb >= g.n_slots) nb = NONE;
                }
                const float d = blk_dist32<CPL>(g, S, T, Q, qn, &B, nb, true, fin, evals);
                int moved = 0;
                if (warp == 0) {
                    float bd = nb != NONE ? d : INFINITY; uint32_t bi = nb;
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) {
                        const float od = __shfl_xor_sync(0xffffffffu, bd, o);
                        const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
                        if (pair_less(od, oi, bd, bi)) { bd = od; bi = oi; }
                    }
                    if (bi != NONE && bd < dcur) { cur = bi; dcur = bd; moved = 1; }
                    if (lane == 0) B.n = moved;     // published by the next barrier below
                }
                __syncthreads();
                moved = B.n;
                __syncthreads();
                if (!moved) break;
            }
        }
        RList<NL> L;
        if (warp == 0) rbeam_seed(S, L, (int)ef, cur, dcur);
        for (;;) {
            uint32_t nb = NONE; bool go = false;
            if (warp == 0) {
                float dc; const uint32_t c = L.pop(dc);
                go = c != NONE;
                if (go) {
                    ++exps;
                    int width; const uint32_t* row = level_row(g, 0, c, width);
                    nb = ((int)lane < width) ? row[lane] : NONE;
                    if (nb != NONE && nb >= g.n_slots) nb = NONE;
                    if (nb != NONE && visited_ts(S.hs, S.hmask, S.tag16, nb)) nb = NONE;
                    __syncwarp();
                }
            }
            const float d = blk_dist32<CPL>(g, S, T, Q, qn, &B, nb, go, fin, evals);
            if (fin) break;
            if (warp == 0) {
                const bool good = nb != NONE && d < L.worst;
                if (good) { int w2; prefetch_l2(level_row(g, 0, nb, w2)); }
                unsigned acc = __ballot_sync(0xffffffffu, good);
                while (acc) {
                    const int src = __ffs(acc) - 1; acc &= acc - 1;
                    const float nd = __shfl_sync(0xffffffffu, d, src);
                    const uint32_t nid = __shfl_sync(0xffffffffu, nb, src);
                    if (nd < L.worst) L.insert(nd, nid);
                }
            }
        }
        if (warp == 0) {
            nout = L.emit_live(g, k, ko, dd);
        }
    }
    if (warp == 0) {
        for (uint32_t i = nout + lane; i < k; i += 32) { ko[i] = 0xFFFFFFFFFFFFFFFFull; dd[i] = INFINITY; }
        __syncwarp();
        fan_out(fan, q, k, ko, dd);
        if (counters && lane == 0) {
            atomicAdd(&counters[0], 1ull);
            atomicAdd(&counters[1], (unsigned long long)evals);
            atomicAdd(&counters[2], (unsigned long long)exps);
        }
    }
}

// ------------------------------------------------- K4a: insertion search on the frozen graph --
// One warp per batch item: greedy descent from the top, then an ef_add beam on every level the
// item lives on; each level's sorted candidate list goes to scratch (cand_id / cand_d).
// Level-l lists of the items with slot % M^l == 0 are stored at item index (slot/M^l - first_row[l]).
__global__ void __launch_bounds__(128) k_build_search(GraphView g, BuildBatch b, uint32_t ef, uint32_t efp, uint32_t hsize,
                                                      unsigned long long* counters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned wpb = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = warp_smem_bytes(g.dpad, efp, hsize);
    WarpSmem S = carve_warp_smem(smem + warp * per_warp, g.dpad, ef, efp, hsize);
    uint32_t evals = 0, exps = 0;
    for (uint32_t it = blockIdx.x * wpb + warp; it < b.count; it += gridDim.x * wpb) {
        const uint32_t slot = b.first_slot + it;
        int lvl = 0;
        while (lvl + 1 <= b.new_top && slot % g.stride[lvl + 1] == 0) ++lvl;
        // mark "no candidates" for every level first
        for (int l = 0; l <= lvl; ++l) {
            const uint32_t idx = slot / g.stride[l] - b.first_row[l];
            if (lane == 0) b.cand_n[l][idx] = 0;
        }
        if (g.n_slots == 0) continue;
        const float qn = warp_load_row(g, S.q4, slot);
        uint32_t cur = 0; float dcur = warp_entry_dist(g, S, qn, 0u, evals);
        for (int l = g.top_level; l > lvl; --l) greedy_level(g, l, S, qn, cur, dcur, evals, exps);
        beam_seed(S, cur, dcur);
        for (int l = min(lvl, g.top_level); l >= 0; --l) {
            beam_level(g, l, S, qn, evals, exps);
            const uint32_t idx = slot / g.stride[l] - b.first_row[l];
            uint32_t* ci = b.cand_id[l] + (size_t)idx * efp; float* cd = b.cand_d[l] + (size_t)idx * efp;
            uint32_t n = 0;
            for (int i = lane; i < (int)efp; i += 32) {
                const uint32_t id = S.L.li[i];
                ci[i] = id == NONE ? NONE : (id & ~EXPANDED); cd[i] = S.L.ld[i];
                n += (id != NONE);
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
            if (lane == 0) b.cand_n[l][idx] = n;
            if (l > 0) beam_reseed_from_list(S);
        }
    }
    if (counters && lane == 0) { atomicAdd(&counters[3], (unsigned long long)evals); atomicAdd(&counters[4], (unsigned long long)exps); }
}

// Register-list variant of K4a (ef_add <= 32 * NL).
template <int NL>
__global__ void __launch_bounds__(128, 6) k_build_search_r(GraphView g, BuildBatch b, uint32_t ef, uint32_t efp, uint32_t hsize,
                                                           unsigned long long* counters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned wpb = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = warp_smem_bytes(g.dpad, 0, hsize);
    WarpSmem S = carve_warp_smem(smem + warp * per_warp, g.dpad, 0, 0, hsize);
    uint32_t evals = 0, exps = 0;
    for (uint32_t it = blockIdx.x * wpb + warp; it < b.count; it += gridDim.x * wpb) {
        const uint32_t slot = b.first_slot + it;
        int lvl = 0;
        while (lvl + 1 <= b.new_top && slot % g.stride[lvl + 1] == 0) ++lvl;
        for (int l = 0; l <= lvl; ++l) {
            const uint32_t idx = slot / g.stride[l] - b.first_row[l];
            if (lane == 0) b.cand_n[l][idx] = 0;
        }
        if (g.n_slots == 0) continue;
        const float qn = warp_load_row(g, S.q4, slot);
        uint32_t cur = 0; float dcur = warp_entry_dist(g, S, qn, 0u, evals);
        for (int l = g.top_level; l > lvl; --l) greedy_level(g, l, S, qn, cur, dcur, evals, exps);
        RList<NL> L;
        rbeam_seed(S, L, (int)ef, cur, dcur);
        for (int l = min(lvl, g.top_level); l >= 0; --l) {
            rbeam_level(g, l, S, L, qn, evals, exps);
            const uint32_t idx = slot / g.stride[l] - b.first_row[l];
            uint32_t* ci = b.cand_id[l] + (size_t)idx * efp; float* cd = b.cand_d[l] + (size_t)idx * efp;
            const uint32_t n = L.store_sorted(ci, cd, (int)efp);
            if (lane == 0) b.cand_n[l][idx] = n;
            if (l > 0) rbeam_reseed(S, L);
        }
    }
    if (counters && lane == 0) { atomicAdd(&counters[3], (unsigned long long)evals); atomicAdd(&counters[4], (unsigned long long)exps); }
}

// Bulk-copy staged variant of K4a (tbeam.cuh).
template <int NL, int CPL>
__global__ void __launch_bounds__(128, VSANN_T_MINB) k_build_search_t(const __grid_constant__ CUtensorMap vmap, int use_vmap, GraphView g, BuildBatch b, uint32_t ef, uint32_t efp, uint32_t hsize, uint32_t stage_rows,
                                                           unsigned long long* counters, uint32_t* work) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = tsmem_bytes(g.dpad, hsize, stage_rows);
    WarpSmem S; WarpStage T;
    carve_tsmem(smem + warp * per_warp, g.dpad, hsize, stage_rows, S, T);
    if (use_vmap) T.tmap = &vmap;
    uint32_t evals = 0, exps = 0;
    for (;;) {
        uint32_t it = 0;
        if (lane == 0) it = atomicAdd(work, 1u);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= b.count) break;
        const uint32_t slot = b.first_slot + it;
        int lvl = 0;
        while (lvl + 1 <= b.new_top && slot % g.stride[lvl + 1] == 0) ++lvl;
        for (int l = 0; l <= lvl; ++l) {
            const uint32_t idx = slot / g.stride[l] - b.first_row[l];
            if (lane == 0) b.cand_n[l][idx] = 0;
        }
        if (g.n_slots == 0) continue;
        const float qn = warp_load_row(g, S.q4, slot);
        QRegs<CPL> Q; Q.load(S.q4);
        uint32_t cur = 0; float dcur = stage_entry_dist<CPL>(g, S, T, Q, qn, 0u, evals);
        for (int l = g.top_level; l > lvl; --l) tgreedy_level<CPL>(g, l, S, T, Q, qn, cur, dcur, evals, exps);
        RList<NL> L;
        rbeam_seed(S, L, (int)ef, cur, dcur);
        for (int l = min(lvl, g.top_level); l >= 0; --l) {
            tbeam_level<NL, CPL>(g, l, S, T, Q, L, qn, evals, exps);
            const uint32_t idx = slot / g.stride[l] - b.first_row[l];
            uint32_t* ci = b.cand_id[l] + (size_t)idx * efp; float* cd = b.cand_d[l] + (size_t)idx * efp;
            const uint32_t n = L.store_sorted(ci, cd, (int)efp);
            if (lane == 0) b.cand_n[l][idx] = n;
            if (l > 0) rbeam_reseed(S, L);
        }
    }
    if (counters && lane == 0) { atomicAdd(&counters[3], (unsigned long long)evals); atomicAdd(&counters[4], (unsigned long long)exps); }
}

// ------------------------------------------------- heuristic selection (shared by K4b / K5) ---
// Candidates arrive ascending by distance to the base through next(); a candidate is kept iff
// it is at least as close to the base as to every neighbour kept so far. The kept ids live one
// per lane. Uses S.q4 as the staging buffer for the candidate row.
struct KeepSet { uint32_t id; int n; };   // lane j < n holds kept id j

__device__ __forceinline__ bool heuristic_try(const GraphView& g, WarpSmem& S, KeepSet& K, uint32_t c, float d_base, uint32_t& evals) {
    bool good = true;
    if (K.n > 0) {
        const float cn = warp_load_row(g, S.q4, c);
        const float dk = warp_dist32(g, S.q4, cn, (int)lane_id() < K.n ? K.id : NONE, S.scr_id, S.scr_d, evals);
        good = __ballot_sync(0xffffffffu, dk < d_base) == 0u;
    }
    if (good) { if ((int)lane_id() == K.n) K.id = c; ++K.n; }
    return good;
}

// The same test with the kept rows resident in shared memory (K4b): the candidate row `cq` (shared, landed by a bulk
// copy) against kept rows 0..kn-1, four at a time (one per 8-lane team), stopping at the first group that holds a
// closer kept neighbour. Same operand order and summation order as warp_dist32 => the same bits, the same decisions.
template <int METRIC>
__device__ __forceinline__ bool kept_any_closer_m(const GraphView& g, const float4* cq, float cn, const float4* kept, int kn,
                                                  float ksq, float d_base, uint32_t& evals) {
    const unsigned lane = lane_id(), team = lane >> 3, tl = lane & 7u;
    const int chunks = g.chunks;
    for (int jb = 0; jb < kn; jb += 4) {   // warp-uniform trip count
        const int j = jb + (int)team; const bool on = j < kn;
        const float4* kr = kept + (size_t)(on ? j : 0) * chunks;
        float2 acc2 = make_float2(0.f, 0.f);
#pragma unroll 4
        for (int c = tl; c < chunks; c += 8) acc_chunk<METRIC>(cq[c], kr[c], acc2);
        float acc = acc2.x + acc2.y;
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        const float xn = METRIC == COS ? __shfl_sync(0xffffffffu, ksq, on ? j : 0) : 0.f;
        const float dk = finish_metric(METRIC, acc, cn, xn);
        evals += (uint32_t)min(4, kn - jb);
        if (__ballot_sync(0xffffffffu, on && dk < d_base)) return true;
    }
    return false;
}
__device__ __forceinline__ bool kept_any_closer(const GraphView& g, const float4* cq, float cn, const float4* kept, int kn,
                                                float ksq, float d_base, uint32_t& evals) {
    switch (g.metric) {
        case L2SQ: return kept_any_closer_m<L2SQ>(g, cq, cn, kept, kn, ksq, d_base, evals);
        case IP: return kept_any_closer_m<IP>(g, cq, cn, kept, kn, ksq, d_base, evals);
        default: return kept_any_closer_m<COS>(g, cq, cn, kept, kn, ksq, d_base, evals);
    }
}
// per-warp extra shared memory of the resident-kept-rows path: 2 mbarriers | 2 candidate rows | M kept rows
__host__ __device__ inline size_t select_extra_bytes(uint32_t dpad, uint32_t M) { return 128 + (size_t)(M + 2) * dpad * 4; }

// ------------------------------------------------- K4b: forward links + reverse-edge inbox ----
// One warp per (item, level). Merges the graph candidates with the intra-batch kNN candidates
// (both ascending), runs the heuristic with cap M, writes the item's row, and registers each
// reverse edge (kept -> item) with the inbox of the kept node's row.
__global__ void __launch_bounds__(128) k_select(GraphView g, BuildBatch b, LinkTables t, uint32_t efp, uint32_t kbp,
                                                unsigned long long* counters, int kept_smem) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned wpb = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = warp_smem_bytes(g.dpad, 32, 32);
    WarpSmem S = carve_warp_smem(smem + warp * per_warp, g.dpad, 32, 32, 32);
    uint32_t evals = 0;
    // resident-kept-rows path: candidate rows arrive by bulk copy (the next one while the current one is tested)
    const uint32_t row_bytes = g.dpad * 4u;
    unsigned char* xb = smem + ((wpb * per_warp + 127) & ~(size_t)127) + warp * select_extra_bytes(g.dpad, g.M);
    const uint32_t bar_s = smem_u32(xb), cand_s = bar_s + 128u;
    const float4* cand0 = reinterpret_cast<const float4*>(xb + 128);   // candidate buffer p starts at cand0 + p * chunks
    float4* kept = reinterpret_cast<float4*>(xb + 128 + 2 * (size_t)row_bytes);
    uint32_t par = 0;
    if (kept_smem) {
        if (lane < 2) mbar_init(bar_s + 8u * lane, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        __syncwarp();
    }
    for (uint32_t w = blockIdx.x * wpb + warp; w < b.total_items; w += gridDim.x * wpb) {
        int l = 0; while (l + 1 <= b.new_top && w >= b.item_off[l + 1]) ++l;
        const uint32_t idx = w - b.item_off[l];
        const uint32_t slot = (b.first_row[l] + idx) * g.stride[l];
        const uint32_t* gi = b.cand_id[l] + (size_t)idx * efp; const float* gd = b.cand_d[l] + (size_t)idx * efp;
        const uint32_t* bi = b.knn_id[l] + (size_t)idx * kbp; const float* bd = b.knn_d[l] + (size_t)idx * kbp;
        const uint32_t gn = b.cand_n[l][idx];
        uint32_t bn = 0;
        if (b.knn_id[l]) { for (uint32_t i = lane; i < kbp; i += 32) bn += (bi[i] != NONE); for (int o = 16; o >= 1; o >>= 1) bn += __shfl_xor_sync(0xffffffffu, bn, o); }
        KeepSet K{NONE, 0};
        uint32_t a = 0, c = 0;
        const int cap = (int)g.M;
        auto next_cand = [&](uint32_t& id, float& d) -> bool {   // merge of the two ascending candidate lists, skipping the item itself
            for (;;) {
                if (!(a < gn || c < bn)) return false;
                const bool take_g = a < gn && (c >= bn || pair_less(gd[a], gi[a], bd[c], bi[c]));
                if (take_g) { id = gi[a]; d = gd[a]; ++a; } else { id = bi[c]; d = bd[c]; ++c; }
                if (id != slot) return true;
            }
        };
        if (kept_smem) {
            auto fetch = [&](uint32_t id, int p) {
                if (lane == 0) {
                    mbar_arrive_expect_tx(bar_s + 8u * p, row_bytes);
                    bulk_g2s(cand_s + (uint32_t)p * row_bytes, g.vecs + (size_t)id * g.dpad, row_bytes, bar_s + 8u * p);
                }
            };
            uint32_t cur_id = NONE, nxt_id = NONE; float cur_d = 0.f, nxt_d = 0.f, ksq = 0.f; int p = 0;
            bool have = next_cand(cur_id, cur_d);
            if (have) fetch(cur_id, p);
            while (have && K.n < cap) {
                const bool have_n = next_cand(nxt_id, nxt_d);
                if (have_n) fetch(nxt_id, p ^ 1);
                const float cn = g.metric == COS ? g.sqn[cur_id] : 0.f;
                mbar_wait(bar_s + 8u * p, (par >> p) & 1u); par ^= 1u << p;
                const bool bad = K.n > 0 && kept_any_closer(g, cand0 + (size_t)p * g.chunks, cn, kept, K.n, ksq, cur_d, evals);
                if (!bad) {
                    float4* dst = kept + (size_t)K.n * g.chunks;
                    for (int ch = lane; ch < (int)g.chunks; ch += 32) dst[ch] = cand0[(size_t)p * g.chunks + ch];
                    if ((int)lane == K.n) { K.id = cur_id; ksq = cn; }
                    ++K.n;
                }
                __syncwarp();
                cur_id = nxt_id; cur_d = nxt_d; have = have_n; p ^= 1;
            }
            if (have) { mbar_wait(bar_s + 8u * p, (par >> p) & 1u); par ^= 1u << p; }   // the prefetched row nobody needs
            __syncwarp();
        } else {
            uint32_t id; float d;
            while (K.n < cap && next_cand(id, d)) heuristic_try(g, S, K, id, d, evals);
        }
        // forward row
        int width; uint32_t* row = const_cast<uint32_t*>(level_row(g, l, slot, width));
        for (int i = lane; i < width; i += 32) row[i] = NONE;
        __syncwarp();
        if ((int)lane < K.n) {
            row[lane] = K.id;
            // reverse edge: K.id <- slot
            const uint32_t r = t.row_base[l] + K.id / g.stride[l];
            const uint32_t e = atomicAdd(t.n_edges, 1u);
            t.edge_row[e] = r; t.edge_src[e] = slot;
            if (atomicAdd(&t.inbox_cnt[r], 1u) == 0u) {
                const uint32_t ti = atomicAdd(t.n_touched, 1u);
                t.touched_row[ti] = r; t.touched_slot[ti] = K.id; t.touched_level[ti] = (uint8_t)l;
                t.inbox_idx[r] = ti;
            }
        }
        __syncwarp();
    }
    if (counters && lane == 0) atomicAdd(&counters[3], (unsigned long long)evals);
}

// Every touched row gets a segment of `inbox` as long as the number of reverse edges it received (counted by k_select),
// so no edge is ever dropped and the merged row is a function of the edge SET: the build is deterministic whatever the
// arrival order (a fixed 32-entry inbox kept "the first 32 to arrive" and made hub rows differ from run to run).
__global__ void k_inbox_alloc(LinkTables t) {
    const uint32_t n = *t.n_touched;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        t.inbox_off[i] = atomicAdd(t.inbox_cursor, t.inbox_cnt[t.touched_row[i]]);
}
__global__ void k_inbox_fill(LinkTables t) {
    const uint32_t n = *t.n_edges;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t ti = t.inbox_idx[t.edge_row[e]];
        const uint32_t pos = atomicAdd(&t.inbox_fill[ti], 1u);
        t.inbox[t.inbox_off[ti] + pos] = t.edge_src[e];
    }
}

// ------------------------------------------------- K5: reverse-link merge / re-prune ----------
// One warp per touched row: candidates = current row + inbox; if they fit the row they are
// appended (sorted by distance), otherwise the heuristic re-prunes to the row cap.
__global__ void __launch_bounds__(128) k_reverse_merge(GraphView g, LinkTables t, unsigned long long* counters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned wpb = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = lane_id();
    const size_t per_warp = warp_smem_bytes(g.dpad, 64, 32);
    WarpSmem S = carve_warp_smem(smem + warp * per_warp, g.dpad, 64, 64, 32);
    const uint32_t ntouched = *t.n_touched;
    uint32_t evals = 0;
    for (uint32_t w = blockIdx.x * wpb + warp; w < ntouched; w += gridDim.x * wpb) {
        const uint32_t v = t.touched_slot[w]; const int l = t.touched_level[w]; const uint32_t r = t.touched_row[w];
        int width; uint32_t* row = const_cast<uint32_t*>(level_row(g, l, v, width));
        const float vn = warp_load_row(g, S.q4, v);
        slist_init(S.L);
        uint32_t total = 0;
        {   // existing links
            uint32_t id = (int)lane < width ? row[lane] : NONE;
            total += __popc(__ballot_sync(0xffffffffu, id != NONE));
            const float d = warp_dist32(g, S.q4, vn, id, S.scr_id, S.scr_d, evals);
            slist_merge(S.L, d, id, S.scr_id, S.scr_d);
        }
        {   // inbox: every reverse edge of this batch, 32 at a time (the list keeps the best 64 of row + inbox by (distance, id))
            const uint32_t nin = t.inbox_cnt[r], off = t.inbox_off[w];
            for (uint32_t c0 = 0; c0 < nin; c0 += 32) {
                uint32_t id = c0 + lane < nin ? t.inbox[off + c0 + lane] : NONE;
                if (id == v) id = NONE;
                const float d = warp_dist32(g, S.q4, vn, id, S.scr_id, S.scr_d, evals);
                slist_merge(S.L, d, id, S.scr_id, S.scr_d);
            }
        }
        // count distinct candidates now in the list
        uint32_t n = 0;
        for (int i = lane; i < 64; i += 32) n += (S.L.li[i] != NONE);
        for (int o = 16; o >= 1; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
        if ((int)n <= width) {
            for (int i = lane; i < width; i += 32) row[i] = i < (int)n ? S.L.li[i] : NONE;
        } else {
            KeepSet K{NONE, 0};
            for (uint32_t i = 0; i < n && K.n < width; ++i) heuristic_try(g, S, K, S.L.li[i], S.L.ld[i], evals);
            // heuristic_try clobbers S.q4 (candidate row staging); the list survives in S.L
            for (int i = lane; i < width; i += 32) row[i] = NONE;
            __syncwarp();
            if ((int)lane < K.n) row[lane] = K.id;
        }
        if (lane == 0) { t.inbox_cnt[r] = 0; t.inbox_fill[w] = 0; }
        __syncwarp();
    }
    if (counters && lane == 0) atomicAdd(&counters[3], (unsigned long long)evals);
}

// ------------------------------------------------- small utility kernels ----------------------
__global__ void k_row_norms(const float* __restrict__ vecs, float* __restrict__ sqn, uint32_t first, uint32_t n, uint32_t dpad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* v = vecs + (size_t)(first + i) * dpad;
    float acc = 0.f;
    for (uint32_t c = 0; c < dpad; ++c) acc = fmaf(v[c], v[c], acc);   // sequential chain = oracle's dot_seq(v, v)
    sqn[first + i] = acc;
}

// src rows (dims floats, dense) -> padded table rows
__global__ void k_scatter_rows(const float* __restrict__ src, float* __restrict__ vecs, uint32_t first, uint32_t n, uint32_t dims, uint32_t dpad) {
    const size_t total = (size_t)n * dpad;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(i / dpad), c = (uint32_t)(i % dpad);
        vecs[(size_t)(first + r) * dpad + c] = c < dims ? src[(size_t)r * dims + c] : 0.f;
    }
}

__global__ void k_tombstone(uint32_t* deleted, const uint32_t* __restrict__ slots, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicOr(&deleted[slots[i] >> 5], 1u << (slots[i] & 31u));
}

// ------------------------------------------------- K6: sharded top-k merge --------------------
// One warp per query; `parts` ascending lists of k (dist, key) each -> global top-k by (dist, key).
__global__ void k_merge_topk(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* __restrict__ keys_in, const float* __restrict__ dist_in,
                             uint64_t* __restrict__ keys_out, float* __restrict__ dist_out) {
    const uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = lane_id();
    if (q >= nq) return;
    // lane p (< 32) owns the head pointer of parts p, p+32, ...; k-step tournament
    uint32_t head[4] = {0, 0, 0, 0};   // supports up to 128 parts
    for (uint32_t o = 0; o < k; ++o) {
        float bd = INFINITY; uint64_t bk = 0xFFFFFFFFFFFFFFFFull; uint32_t bp = NONE;
        for (uint32_t j = 0, p = lane; p < parts && j < 4; ++j, p += 32) {
            if (head[j] < k) {
                const size_t off = ((size_t)p * nq + q) * k + head[j];
                const float d = dist_in[off]; const uint64_t key = keys_in[off];
                if (d < bd || (d == bd && key < bk)) { bd = d; bk = key; bp = p; }
            }
        }
        float wd = bd; uint64_t wk = bk; uint32_t wp = bp;
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, wd, s);
            const uint64_t ok = __shfl_xor_sync(0xffffffffu, wk, s);
            const uint32_t op = __shfl_xor_sync(0xffffffffu, wp, s);
            if (od < wd || (od == wd && (ok < wk || (ok == wk && op < wp)))) { wd = od; wk = ok; wp = op; }
        }
        if (lane == 0) { keys_out[(size_t)q * k + o] = wk; dist_out[(size_t)q * k + o] = wd; }
        if (wp != NONE && (wp & 31u) == lane) ++head[wp >> 5];
    }
}

// Fused gather + merge: part p's lists are read where they are - part_keys[p] / part_dist[p] may point
// into a PEER GPU's memory mapped over NVLink / NVSwitch (symmetric memory) - so the exchange step needs no all-gather and
// no gathered copy in HBM. A dependent tournament over remote loads would pay the NVLink latency k times, so the warp first
// pulls all parts x k candidates of its query into shared memory in one round of independent loads, then merges locally.
// STAGED = false: lists too large for shared memory are read in place (same tournament as k_merge_topk).
// flags != null: the cross-rank barrier is part of this kernel. Rank `my_part` publishes "my lists of step `epoch` are complete"
// (they are: this kernel follows the rank's traversal kernel on its stream) with a system-scope release store into its flag word,
// which every peer has mapped, and every block waits (system-scope acquire loads over NVLink) until each peer's flag has reached
// `epoch` before it pulls that peer's lists: one launch after the traversal, no separate barrier kernel, no launch gap in between.
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

// The barrier half of the fused exchange (see below): called by every thread of every block.
__device__ __forceinline__ void merge_gate(uint32_t* const* __restrict__ flags, uint32_t parts, uint32_t my_part, uint32_t epoch) {
    if (blockIdx.x == 0) {
        // ONE block per GPU polls the peers: thousands of pollers on a peer's flag word contend with that peer's traversal
        // kernel for its NVLink ingress (measured at N = 2: end-to-end step 0.80 ms with every block polling, against 0.64 ms).
        // One LANE per peer: a poll is an NVLink round trip, seven of them in a row were most of the barrier at N = 8.
        if (threadIdx.x < 32) {
            uint32_t* mine = flags[my_part];      // word 0: what the peers read; word 1: "every peer has arrived", read by this GPU's other blocks
            if (threadIdx.x == 0) { __threadfence_system(); st_release_sys(mine, epoch); }
            for (uint32_t p = threadIdx.x; p < parts; p += 32) {
                if (p == my_part) continue;
                while ((int32_t)(ld_acquire_sys(flags[p]) - epoch) < 0) __nanosleep(200);
            }
            __syncwarp();
            if (threadIdx.x == 0) { __threadfence(); asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(mine + 1), "r"(epoch) : "memory"); }
        }
    } else if (threadIdx.x == 0) {
        const uint32_t* mine = flags[my_part];
        uint32_t v;
        do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(mine + 1) : "memory"); if ((int32_t)(v - epoch) < 0) __nanosleep(100); } while ((int32_t)(v - epoch) < 0);
    }
    __syncthreads();
}

template <bool STAGED>
__global__ void k_merge_topk_ptrs(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* const* __restrict__ part_keys,
                                  const float* const* __restrict__ part_dist, uint32_t* const* __restrict__ flags, uint32_t my_part, uint32_t epoch,
                                  uint64_t* __restrict__ keys_out, float* __restrict__ dist_out) {
    extern __shared__ __align__(16) unsigned char msm[];
    if (flags) merge_gate(flags, parts, my_part, epoch);
    const uint32_t warp = threadIdx.x >> 5, q = blockIdx.x * (blockDim.x >> 5) + warp, lane = lane_id();
    if (q >= nq) return;
    const uint32_t tot = parts * k;
    uint64_t* sk = reinterpret_cast<uint64_t*>(msm) + (size_t)warp * tot;
    float* sd = reinterpret_cast<float*>(msm + (size_t)(blockDim.x >> 5) * tot * 8) + (size_t)warp * tot;
    if (STAGED) {
        for (uint32_t i = lane; i < tot; i += 32) {
            const uint32_t p = i / k, j = i - p * k;
            sk[i] = part_keys[p][(size_t)q * k + j]; sd[i] = part_dist[p][(size_t)q * k + j];
        }
        __syncwarp();
    }
    uint32_t head[4] = {0, 0, 0, 0};   // lane l owns the heads of parts l, l + 32, ...: up to 128 parts
    for (uint32_t o = 0; o < k; ++o) {
        float bd = INFINITY; uint64_t bk = 0xFFFFFFFFFFFFFFFFull; uint32_t bp = NONE;
        for (uint32_t j = 0, p = lane; p < parts && j < 4; ++j, p += 32) {
            if (head[j] < k) {
                float d; uint64_t key;
                if (STAGED) { d = sd[p * k + head[j]]; key = sk[p * k + head[j]]; }
                else { d = part_dist[p][(size_t)q * k + head[j]]; key = part_keys[p][(size_t)q * k + head[j]]; }
                if (d < bd || (d == bd && key < bk)) { bd = d; bk = key; bp = p; }
            }
        }
        float wd = bd; uint64_t wk = bk; uint32_t wp = bp;
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, wd, s);
            const uint64_t ok = __shfl_xor_sync(0xffffffffu, wk, s);
            const uint32_t op = __shfl_xor_sync(0xffffffffu, wp, s);
            if (od < wd || (od == wd && (ok < wk || (ok == wk && op < wp)))) { wd = od; wk = ok; wp = op; }
        }
        if (lane == 0) { keys_out[(size_t)q * k + o] = wk; dist_out[(size_t)q * k + o] = wd; }
        if (wp != NONE && (wp & 31u) == lane) ++head[wp >> 5];
    }
}

// The same exchange for lists that fit shared memory, QB queries per block: measured on two B200s with every list ready, the
// warp-per-query tournament above took 46 us for 10 000 x 10 lists from 2 parts and 123 us for 4096 x 100 (the flag barrier only
// 8 us of the fused kernel): k dependent rounds of a 32-lane reduction to find ONE output, and a few 8-byte peer loads per lane.
// Here the block pulls the lists of its QB queries from every part with coalesced loads (a part's lists of consecutive queries are
// one contiguous run), and every (query, part, position) element finds its OWN place in the merged order: position + the number
// of elements of the other parts that precede it - a binary search per other part over its sorted list, no rounds, no shuffles.
// Order: (distance, key), then part for full ties (padding entries) - the order of k_merge_topk / merge_topk_host.
__global__ void __launch_bounds__(256) k_merge_topk_rank(uint32_t parts, uint32_t nq, uint32_t k, uint32_t qb, const uint64_t* const* __restrict__ part_keys,
                                                         const float* const* __restrict__ part_dist, uint32_t* const* __restrict__ flags, uint32_t my_part,
                                                         uint32_t epoch, uint64_t* __restrict__ keys_out, float* __restrict__ dist_out) {
    extern __shared__ __align__(16) unsigned char msm[];
    if (flags) merge_gate(flags, parts, my_part, epoch);
    const uint32_t q0 = blockIdx.x * qb;
    if (q0 >= nq) return;
    const uint32_t nqb = min(qb, nq - q0), per_part = nqb * k;
    uint64_t* sk = reinterpret_cast<uint64_t*>(msm);                              // [part][query][k]
    float* sd = reinterpret_cast<float*>(msm + (size_t)parts * qb * k * 8);
    // all parts' loads of an element leave before any of them is stored: one NVLink round trip per block, not one per peer
    for (uint32_t p0 = 0; p0 < parts; p0 += 8) {
        const uint64_t* gk[8]; const float* gd[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { const uint32_t p = min(p0 + u, parts - 1); gk[u] = part_keys[p] + (size_t)q0 * k; gd[u] = part_dist[p] + (size_t)q0 * k; }
        for (uint32_t i = threadIdx.x; i < per_part; i += blockDim.x) {
            uint64_t kk[8]; float dd[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) if (p0 + u < parts) { kk[u] = gk[u][i]; dd[u] = gd[u][i]; }
#pragma unroll
            for (int u = 0; u < 8; ++u) if (p0 + u < parts) { sk[(size_t)(p0 + u) * qb * k + i] = kk[u]; sd[(size_t)(p0 + u) * qb * k + i] = dd[u]; }
        }
    }
    __syncthreads();
    // bound[q] = the smallest LAST distance over the parts: that part alone holds k entries <= bound, so nothing beyond it can be
    // among the k best (with 8 parts seven eighths of the entries leave here, before any search)
    float* bound = reinterpret_cast<float*>(msm + (size_t)parts * qb * k * 12);
    for (uint32_t ql = threadIdx.x; ql < nqb; ql += blockDim.x) {
        float b = INFINITY;
        for (uint32_t p = 0; p < parts; ++p) b = fminf(b, sd[(size_t)p * qb * k + ql * k + (k - 1)]);
        bound[ql] = b;
    }
    __syncthreads();
    const uint32_t total = parts * per_part;
    for (uint32_t e = threadIdx.x; e < total; e += blockDim.x) {
        const uint32_t p = e / per_part, r = e - p * per_part, ql = r / k, j = r - ql * k;
        const float d = sd[(size_t)p * qb * k + r]; const uint64_t key = sk[(size_t)p * qb * k + r];
        if (d > bound[ql]) continue;
        uint32_t rank = j;
        for (uint32_t o = 0; o < parts && rank < k; ++o) {
            if (o == p) continue;
            const float* od = sd + (size_t)o * qb * k + ql * k; const uint64_t* ok = sk + (size_t)o * qb * k + ql * k;
            // number of entries of part o that precede (d, key, p): lower bound for o > p, upper bound for o < p
            uint32_t lo = 0, hi = min(k, k - rank);      // more than k - rank of them would put this element past the output anyway
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                const float md = od[mid]; const uint64_t mk = ok[mid];
                const bool before = md < d || (md == d && (mk < key || (mk == key && o < p)));
                if (before) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        if (rank < k) { keys_out[(size_t)(q0 + ql) * k + rank] = key; dist_out[(size_t)(q0 + ql) * k + rank] = d; }
    }
}

// ------------------------------------------------- launchers ----------------------------------
static inline uint32_t next_pow2(uint32_t v) { uint32_t p = 1; while (p < v) p <<= 1; return p; }

static SearchPlan plan_search(uint32_t dpad, uint32_t ef, int sm_count, size_t smem_limit, uint32_t hash_factor) {
    SearchPlan p;
    p.efp = (ef + 31u) & ~31u;
    uint32_t hs = next_pow2(ef * hash_factor);
    if (hs < 1024) hs = 1024;
    if (hs > 16384) hs = 16384;
    size_t per_warp = warp_smem_bytes(dpad, p.efp, hs);
    while (per_warp > smem_limit && hs > 256) { hs >>= 1; per_warp = warp_smem_bytes(dpad, p.efp, hs); }
    p.hsize = hs;
    uint32_t wpb = 4;
    while (wpb > 1 && per_warp * wpb > smem_limit / 2) wpb >>= 1;
    p.warps_per_block = wpb;
    p.smem = per_warp * wpb;
    p.blocks_per_sm = (uint32_t)(smem_limit / (p.smem + 1024));
    if (p.blocks_per_sm < 1) p.blocks_per_sm = 1;
    if (p.blocks_per_sm * wpb > 48) p.blocks_per_sm = 48 / wpb;
    (void)sm_count;
    return p;
}


// Work counter of one persistent launch: a slot of the index's ring, zeroed in stream order.
static std::atomic<uint32_t> g_work_next{0};
static cudaError_t next_work_counter(const DeviceInfo& dev, cudaStream_t st, uint32_t** out) {
    if (!dev.work_ring) return cudaErrorInvalidValue;
    *out = dev.work_ring + (g_work_next.fetch_add(1) % WORK_RING_SLOTS);
    return cudaMemsetAsync(*out, 0, sizeof(uint32_t), st);
}

// ---- staged (bulk-copy) launch geometry: per-warp staging budget -> rows per round ----
static inline uint32_t stage_rows_for(const GraphView& g, const DeviceInfo& dev) {
    const uint32_t row_bytes = g.dpad * 4;
    uint32_t rows = (dev.stage_bytes / row_bytes) & ~3u;     // whole gather4 groups
    if (rows > 32) rows = 32;
    if (rows < 4) rows = 4;
    return rows;
}
// vecs as a [slots][dpad] tensor with a one-row box for tile::gather4 (rows of at most 256 elements)
static inline int vecs_tmap(const GraphView& g, const DeviceInfo& dev, CUtensorMap* m) {
    memset(m, 0, sizeof(*m));
    if (!dev.gather4 || g.dpad > 256 || (g.dpad & 7u) || g.n_slots == 0) return 0;   // groups of 4 rows must stay 128 B aligned
    return make_tmap_f32(m, g.vecs, g.n_slots, g.dpad, g.dpad, g.dpad, 1, CU_TENSOR_MAP_SWIZZLE_NONE) ? 1 : 0;
}
static inline uint32_t hash_size_for(uint32_t ef, const DeviceInfo& dev, const GraphView& g, uint32_t stage_rows) {
    uint32_t hs = next_pow2(ef * dev.hash_factor);
    if (hs < dev.hash_min) hs = dev.hash_min;
    if (hs > 16384) hs = 16384;
    // 16-bit tags where the table would otherwise be the larger part of the warp's shared memory (common.cuh HASH_TAGS)
    static const int force = [] { const char* e = getenv("VSANN_HASH_TAGS"); return e ? atoi(e) : -1; }();   // A/B: 0 never, 1 whenever > 1024 slots
    const bool tags = hs > 1024 && (force == 1 || (force != 0 && (size_t)hs * 4 >= (size_t)stage_rows * g.dpad * 4));
    return tags ? (hs | HASH_TAGS) : hs;
}

template <int NL, int CPL>
static cudaError_t launch_search_t(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                                   uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st, const OutFan& fan) {
    const uint32_t rows = stage_rows_for(g, dev), hs = hash_size_for(ef, dev, g, rows);
    uint32_t wpb = 4;
    // wide lists (ef > 256: NL = 16 / 32 list registers per lane) with long rows: fewer warps per block so that hash + staging fit
    while (NL > 8 && wpb > 1 && tsmem_bytes(g.dpad, hs, rows) * wpb > dev.smem_per_block) wpb >>= 1;
    const size_t smem = tsmem_bytes(g.dpad, hs, rows) * wpb;
    if (smem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_search_t<NL, CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    uint32_t* work = nullptr;
    if ((e = next_work_counter(dev, st, &work)) != cudaSuccess) return e;
    const uint32_t blocks = std::min<uint32_t>((nq + wpb - 1) / wpb, (uint32_t)dev.sm_count * (NL <= 8 ? VSANN_T_MINB : 8));
    CUtensorMap vmap; const int use_vmap = vecs_tmap(g, dev, &vmap);
    k_search_t<NL, CPL><<<blocks, wpb * 32, smem, st>>>(vmap, use_vmap, g, d_queries, nq, k, ef, hs, rows, d_keys, d_dist, counters, work, fan);
    return cudaGetLastError();
}
template <int NL>
static cudaError_t launch_search_tc(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                                    uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st, const OutFan& fan) {
    if (g.chunks == 32) return launch_search_t<NL, 4>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
    if (g.chunks == 24) return launch_search_t<NL, 3>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
    return launch_search_t<NL, 0>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
}

template <int NL, int CPL>
static cudaError_t launch_build_search_t(const GraphView& g, const BuildBatch& b, uint32_t ef, const DeviceInfo& dev,
                                         unsigned long long* counters, cudaStream_t st) {
    const uint32_t rows = stage_rows_for(g, dev), hs = hash_size_for(ef, dev, g, rows), wpb = 4;
    const size_t smem = tsmem_bytes(g.dpad, hs, rows) * wpb;
    if (smem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_build_search_t<NL, CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    uint32_t* work = nullptr;
    if ((e = next_work_counter(dev, st, &work)) != cudaSuccess) return e;
    const uint32_t blocks = std::min<uint32_t>((b.count + wpb - 1) / wpb, (uint32_t)dev.sm_count * VSANN_T_MINB);
    CUtensorMap vmap; const int use_vmap = vecs_tmap(g, dev, &vmap);
    k_build_search_t<NL, CPL><<<blocks, wpb * 32, smem, st>>>(vmap, use_vmap, g, b, ef, build_efp(ef), hs, rows, counters, work);
    return cudaGetLastError();
}
template <int NL>
static cudaError_t launch_build_search_tc(const GraphView& g, const BuildBatch& b, uint32_t ef, const DeviceInfo& dev,
                                          unsigned long long* counters, cudaStream_t st) {
    if (g.chunks == 32) return launch_build_search_t<NL, 4>(g, b, ef, dev, counters, st);
    if (g.chunks == 24) return launch_build_search_t<NL, 3>(g, b, ef, dev, counters, st);
    return launch_build_search_t<NL, 0>(g, b, ef, dev, counters, st);
}

template <int NL>
static cudaError_t launch_search_r(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                                   uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st) {
    uint32_t hs = next_pow2(ef * dev.hash_factor);
    if (hs < 1024) hs = 1024;
    if (hs > 16384) hs = 16384;
    const uint32_t wpb = 4;
    const size_t smem = warp_smem_bytes(g.dpad, 0, hs) * wpb;
    if (smem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_search_r<NL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_search_r<NL><<<(nq + wpb - 1) / wpb, wpb * 32, smem, st>>>(g, d_queries, nq, k, ef, hs, d_keys, d_dist, counters);
    return cudaGetLastError();
}

template <int NL, int CPL>
static cudaError_t launch_search_blk_t(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                                       uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st, const OutFan& fan) {
    const uint32_t rows = stage_rows_for(g, dev), hs = hash_size_for(ef, dev, g, rows);
    const size_t per_warp = tsmem_bytes(g.dpad, hs, rows);
    uint32_t W = 8;
    while (W > 1 && per_warp * W + 1024 > dev.smem_per_block) --W;
    if (W < 2) return cudaErrorInvalidConfiguration;
    const size_t smem = per_warp * W;
    cudaError_t e = cudaFuncSetAttribute(k_search_blk<NL, CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    CUtensorMap vmap; const int use_vmap = vecs_tmap(g, dev, &vmap);
    k_search_blk<NL, CPL><<<nq, W * 32, smem, st>>>(vmap, use_vmap, g, d_queries, nq, k, ef, hs, rows, d_keys, d_dist, counters, fan);
    return cudaGetLastError();
}
template <int NL>
static cudaError_t launch_search_blk_c(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                                       uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st, const OutFan& fan) {
    if (g.chunks == 32) return launch_search_blk_t<NL, 4>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
    return launch_search_blk_t<NL, 0>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
}

cudaError_t launch_search(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                          uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st, const OutFan* fan_in) {
    const OutFan fan = fan_in ? *fan_in : OutFan{};
    if (nq == 0) return cudaSuccess;
    // small batches: a block per query (K3b) - the SMs would idle otherwise and each query is latency bound on one warp
    static const uint32_t blk_max = [] { const char* e = getenv("VSANN_BLOCK_QUERY_MAX"); return e ? (uint32_t)atoi(e) : 0xFFFFFFFFu; }();
    const uint32_t blk_limit = blk_max != 0xFFFFFFFFu ? blk_max : 2u * (uint32_t)dev.sm_count;
    if (dev.stage_bytes && !dev.force_smem_list && ef <= 256 && nq <= blk_limit) {
        cudaError_t e;
        if (ef <= 32) e = launch_search_blk_c<1>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else if (ef <= 64) e = launch_search_blk_c<2>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else if (ef <= 128) e = launch_search_blk_c<4>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else e = launch_search_blk_c<8>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        if (e != cudaErrorInvalidConfiguration) return e;
    }
    if (dev.stage_bytes && !dev.force_smem_list && ef <= 1024) {
        cudaError_t e;
        if (ef <= 32) e = launch_search_tc<1>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else if (ef <= 64) e = launch_search_tc<2>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else if (ef <= 128) e = launch_search_tc<4>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else if (ef <= 256) e = launch_search_tc<8>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        // ef up to 1024 (configs[4]'s recall sweep): the register list still beats the shared-memory list kernel
        // (5M x 1536, k = 100: ef = 512 took 67 ms per 4096 queries there against 17.6 ms at ef = 256)
        else if (ef <= 512) e = launch_search_tc<16>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        else e = launch_search_tc<32>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st, fan);
        if (e != cudaErrorInvalidConfiguration) return e;   // rows too long for the staging area: register-gather kernels below
    }
    if (fan.n) return cudaErrorNotSupported;       // the register-gather and shared-memory-list kernels write one result buffer only
    if (!dev.force_smem_list) {
        if (ef <= 32) return launch_search_r<1>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st);
        if (ef <= 64) return launch_search_r<2>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st);
        if (ef <= 128) return launch_search_r<4>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st);
        if (ef <= 256) return launch_search_r<8>(g, d_queries, nq, k, ef, dev, d_keys, d_dist, counters, st);
    }
    const SearchPlan p = plan_search(g.dpad, ef, dev.sm_count, dev.smem_per_block, dev.hash_factor);
    if (p.smem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem);
    if (e != cudaSuccess) return e;
    const uint32_t blocks = (nq + p.warps_per_block - 1) / p.warps_per_block;
    k_search<<<blocks, p.warps_per_block * 32, p.smem, st>>>(g, d_queries, nq, k, ef, p.efp, p.hsize, d_keys, d_dist, counters);
    return cudaGetLastError();
}

uint32_t build_efp(uint32_t ef) { return (ef + 31u) & ~31u; }

template <int NL>
static cudaError_t launch_build_search_r(const GraphView& g, const BuildBatch& b, uint32_t ef, const DeviceInfo& dev,
                                         unsigned long long* counters, cudaStream_t st) {
    uint32_t hs = next_pow2(ef * dev.hash_factor);
    if (hs < 1024) hs = 1024;
    if (hs > 16384) hs = 16384;
    const uint32_t wpb = 4;
    const size_t smem = warp_smem_bytes(g.dpad, 0, hs) * wpb;
    if (smem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_build_search_r<NL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_build_search_r<NL><<<(b.count + wpb - 1) / wpb, wpb * 32, smem, st>>>(g, b, ef, build_efp(ef), hs, counters);
    return cudaGetLastError();
}

cudaError_t launch_build_search(const GraphView& g, const BuildBatch& b, uint32_t ef, const DeviceInfo& dev,
                                unsigned long long* counters, cudaStream_t st) {
    if (b.count == 0) return cudaSuccess;
    if (dev.stage_bytes && !dev.force_smem_list && ef <= 256) {
        cudaError_t e;
        if (ef <= 32) e = launch_build_search_tc<1>(g, b, ef, dev, counters, st);
        else if (ef <= 64) e = launch_build_search_tc<2>(g, b, ef, dev, counters, st);
        else if (ef <= 128) e = launch_build_search_tc<4>(g, b, ef, dev, counters, st);
        else e = launch_build_search_tc<8>(g, b, ef, dev, counters, st);
        if (e != cudaErrorInvalidConfiguration) return e;
    }
    if (!dev.force_smem_list) {
        if (ef <= 32) return launch_build_search_r<1>(g, b, ef, dev, counters, st);
        if (ef <= 64) return launch_build_search_r<2>(g, b, ef, dev, counters, st);
        if (ef <= 128) return launch_build_search_r<4>(g, b, ef, dev, counters, st);
        if (ef <= 256) return launch_build_search_r<8>(g, b, ef, dev, counters, st);
    }
    const SearchPlan p = plan_search(g.dpad, ef, dev.sm_count, dev.smem_per_block, dev.hash_factor);
    if (p.smem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_build_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem);
    if (e != cudaSuccess) return e;
    const uint32_t blocks = (b.count + p.warps_per_block - 1) / p.warps_per_block;
    k_build_search<<<blocks, p.warps_per_block * 32, p.smem, st>>>(g, b, ef, p.efp, p.hsize, counters);
    return cudaGetLastError();
}

cudaError_t launch_link(const GraphView& g, const BuildBatch& b, const LinkTables& t, uint32_t efp, uint32_t kbp, const DeviceInfo& dev,
                        unsigned long long* counters, cudaStream_t st) {
    if (b.total_items == 0) return cudaSuccess;
    const uint32_t wpb = 4;
    size_t smem = warp_smem_bytes(g.dpad, 32, 32) * wpb;
    // kept rows resident in shared memory while they fit a 24 KB per-warp budget (d <= 340 at M = 16); else gathers from L1/L2
    const int kept_smem = dev.select_smem && select_extra_bytes(g.dpad, g.M) <= 24 * 1024;
    if (kept_smem) smem = ((smem + 127) & ~(size_t)127) + wpb * select_extra_bytes(g.dpad, g.M);
    cudaError_t e = cudaFuncSetAttribute(k_select, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    uint32_t blocks = (b.total_items + wpb - 1) / wpb;
    uint32_t maxb = dev.sm_count * 8; if (blocks > maxb) blocks = maxb;
    k_select<<<blocks, wpb * 32, smem, st>>>(g, b, t, efp, kbp, counters, kept_smem);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    k_inbox_alloc<<<dev.sm_count * 4, 256, 0, st>>>(t);
    k_inbox_fill<<<dev.sm_count * 4, 256, 0, st>>>(t);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    smem = warp_smem_bytes(g.dpad, 64, 32) * wpb;
    e = cudaFuncSetAttribute(k_reverse_merge, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    // touched rows <= reverse edges <= total_items * M (device-side count bounds the loop)
    uint64_t maxw = (uint64_t)b.total_items * g.M;
    blocks = (uint32_t)((maxw + wpb - 1) / wpb); if (blocks > maxb) blocks = maxb;
    k_reverse_merge<<<blocks, wpb * 32, smem, st>>>(g, t, counters);
    return cudaGetLastError();
}

cudaError_t launch_row_prepare(const float* d_src, float* vecs, float* sqn, uint32_t first, uint32_t n, uint32_t dims, uint32_t dpad, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const size_t total = (size_t)n * dpad;
    uint32_t blocks = (uint32_t)((total + 255) / 256); if (blocks > 148 * 16) blocks = 148 * 16;
    if (d_src) k_scatter_rows<<<blocks, 256, 0, st>>>(d_src, vecs, first, n, dims, dpad);   // null: the rows are already in the table (dims == dpad, copied straight from the host)
    k_row_norms<<<(n + 127) / 128, 128, 0, st>>>(vecs, sqn, first, n, dpad);
    return cudaGetLastError();
}

cudaError_t launch_tombstone(uint32_t* deleted, const uint32_t* d_slots, uint32_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_tombstone<<<(n + 255) / 256, 256, 0, st>>>(deleted, d_slots, n);
    return cudaGetLastError();
}

// found[q] = number of leading real keys of query q's result row (rows are padded with UINT64_MAX). Counting on the host
// cost 65 us per 10 000 x 10 results (a single-thread pass over 800 KB that had just arrived by DMA): a tenth of the step.
__global__ void k_count_found(const uint64_t* __restrict__ keys, uint32_t nq, uint32_t k, uint32_t* __restrict__ found) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const uint64_t* r = keys + (size_t)q * k;
    uint32_t f = k;
    while (f > 0 && r[f - 1] == 0xFFFFFFFFFFFFFFFFull) --f;
    found[q] = f;
}
cudaError_t launch_count_found(const uint64_t* keys, uint32_t nq, uint32_t k, uint32_t* found, cudaStream_t st) {
    if (nq == 0) return cudaSuccess;
    k_count_found<<<(nq + 255) / 256, 256, 0, st>>>(keys, nq, k, found);
    return cudaGetLastError();
}

cudaError_t launch_merge_topk(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* keys_in, const float* dist_in,
                              uint64_t* keys_out, float* dist_out, cudaStream_t st) {
    if (nq == 0 || k == 0) return cudaSuccess;
    if (parts > 128) return cudaErrorInvalidValue;
    const uint32_t threads = 128, qpb = threads / 32;
    k_merge_topk<<<(nq + qpb - 1) / qpb, threads, 0, st>>>(parts, nq, k, keys_in, dist_in, keys_out, dist_out);
    return cudaGetLastError();
}

cudaError_t launch_merge_topk_flags(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* const* part_keys, const float* const* part_dist,
                                    uint32_t* const* flags, uint32_t my_part, uint32_t epoch, uint64_t* keys_out, float* dist_out, cudaStream_t st) {
    if (k == 0 || parts == 0) return cudaSuccess;
    if (nq == 0 && !flags) return cudaSuccess;
    if (parts > 128) return cudaErrorInvalidValue;
    // lists that fit shared memory: QB queries per block through the rank merge
    const size_t per_query = (size_t)parts * k * 12;
    const char* me = getenv("VSANN_MERGE_RANK");       // 0 = the warp tournament (A/B switch)
    if (per_query <= 64 * 1024 && !(me && atoi(me) == 0)) {
        uint32_t qb = (uint32_t)std::min<size_t>(32, (64 * 1024) / per_query);
        while (qb > 1 && (nq + qb - 1) / qb < 2u * 148u) qb = (qb + 1) / 2;          // small batches: enough blocks to fill the GPU
        const size_t smem = (size_t)qb * per_query + (size_t)qb * 4;
        cudaError_t e = cudaFuncSetAttribute(k_merge_topk_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        const uint32_t blocks = std::max<uint32_t>(1, (nq + qb - 1) / qb);   // at least one block: the flag must be published even for an empty batch
        const uint32_t work = qb * parts * k, threads = work >= 256 ? 256 : (work >= 128 ? 128 : 64);
        k_merge_topk_rank<<<blocks, threads, smem, st>>>(parts, nq, k, qb, part_keys, part_dist, flags, my_part, epoch, keys_out, dist_out);
        return cudaGetLastError();
    }
    const uint32_t threads = 128, qpb = threads / 32;
    const size_t smem = (size_t)qpb * parts * k * 12;
    const uint32_t blocks = std::max<uint32_t>(1, (nq + qpb - 1) / qpb);   // at least one block: the flag must be published even for an empty batch
    if (smem <= 96 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_merge_topk_ptrs<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        k_merge_topk_ptrs<true><<<blocks, threads, smem, st>>>(parts, nq, k, part_keys, part_dist, flags, my_part, epoch, keys_out, dist_out);
    } else {
        k_merge_topk_ptrs<false><<<blocks, threads, 0, st>>>(parts, nq, k, part_keys, part_dist, flags, my_part, epoch, keys_out, dist_out);
    }
    return cudaGetLastError();
}
cudaError_t launch_merge_topk_ptrs(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* const* part_keys, const float* const* part_dist,
                                   uint64_t* keys_out, float* dist_out, cudaStream_t st) {
    return launch_merge_topk_flags(parts, nq, k, part_keys, part_dist, nullptr, 0, 0, keys_out, dist_out, st);
}

}  // namespace vsann
