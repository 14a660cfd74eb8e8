// exact_kernels.cu — K1/K2: exact (brute-force) k-NN, FP32, bit-identical to the oracle.
//
// The north star's "reference's exact search" has no call site in the reference;
// it is defined by oracle/oracle_vsann.cpp exact_search: f32 distance with a sequential fmaf
// chain over the dimensions in index order, results ordered by (distance asc, key asc).
// This kernel keeps that exact accumulation order — every (query, row) pair owns one
// accumulator that is updated for k = 0, 1, ..., d-1 in order — so ids AND distances match the
// oracle bit for bit. It is the register-tiled FP32 form (128 x 128 tile, 8 x 8 per thread);
// the same tiles also produce the intra-batch neighbours of an insertion batch (K4).
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"
#include "tbeam.cuh"
#include "tmap.h"

namespace vsann {

constexpr int BQ = 128, BC = 128, KC = 16, EX_THREADS = 256, CB = 16;   // CB: per-query candidate buffer per round

struct ExactPlan { uint32_t nqt, nchunks, rows_per_chunk, kp; };

static ExactPlan plan_exact(uint32_t nq, uint32_t n, uint32_t k, int sm_count) {
    ExactPlan p;
    p.nqt = (nq + BQ - 1) / BQ;
    p.kp = (k + 31u) & ~31u;
    uint32_t want = (2u * (uint32_t)sm_count + p.nqt - 1) / p.nqt;
    const uint32_t tiles = (n + BC - 1) / BC;
    if (want > tiles) want = tiles;
    if (want < 1) want = 1;
    p.rows_per_chunk = ((tiles + want - 1) / want) * BC;
    p.nchunks = (n + p.rows_per_chunk - 1) / p.rows_per_chunk;
    if (p.nchunks < 1) p.nchunks = 1;
    return p;
}

static uint32_t max_chunks(uint32_t nq, int sm_count) {
    const uint32_t nqt = (nq + BQ - 1) / BQ;
    uint32_t want = (2u * (uint32_t)sm_count + nqt - 1) / nqt;
    return want < 1 ? 1 : want;
}

// scratch layout: qn[nq] | par[nq * maxchunks] (u8) | part_d[nq*maxchunks*2*kp] | part_i[...]
size_t exact_scratch_bytes(uint32_t nq, uint32_t k, int sm_count) {
    const size_t kp = (k + 31u) & ~31u, mc = max_chunks(nq, sm_count);
    size_t b = ((size_t)nq * 4 + 255) & ~(size_t)255;
    b += ((size_t)nq * mc + 255) & ~(size_t)255;
    b += (size_t)nq * mc * 2 * kp * 8 + 512;
    return b;
}

// Upper bound of exact_scratch_bytes over every nq <= max_nq: nq * chunks(nq) <= 256 * sm + nq.
size_t exact_scratch_bytes_upto(uint32_t max_nq, uint32_t k, int sm_count) {
    const size_t kp = (k + 31u) & ~31u, rows = (size_t)256 * sm_count + max_nq;
    return (((size_t)max_nq * 4 + 255) & ~(size_t)255) + ((rows + 255) & ~(size_t)255) + rows * 2 * kp * 8 + 512;
}

__global__ void k_query_norms(ExactQueries q, float* qn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= q.nq) return;
    const float* v = q.q + (size_t)i * q.stride;
    float acc = 0.f;
    for (uint32_t c = 0; c < q.dpad; ++c) acc = fmaf(v[c], v[c], acc);
    qn[i] = acc;
}

template <int METRIC, class Less>
__global__ void __launch_bounds__(EX_THREADS, 2)
k_exact_tiles(ExactCorpus C, ExactQueries Q, ExactPlan P, uint32_t k, const float* __restrict__ qn,
              float* part_d, uint32_t* part_i, uint8_t* par_out, Less less) {
    __shared__ __align__(16) float qs[KC][BQ];
    __shared__ __align__(16) float cs[KC][BC];
    __shared__ float thr[BQ];
    __shared__ uint32_t ccnt[BQ];
    __shared__ float cbuf_d[BQ][CB];
    __shared__ uint32_t cbuf_i[BQ][CB];
    __shared__ uint8_t par[BQ];
    __shared__ uint32_t scr_id[EX_THREADS / 32][32];
    __shared__ float scr_d[EX_THREADS / 32][32];

    const int t = threadIdx.x, tx = t & 15, ty = t >> 4, warp = t >> 5, lane = t & 31;
    const uint32_t chunk = blockIdx.x, q0 = blockIdx.y * BQ;
    const uint32_t r_begin = chunk * P.rows_per_chunk;
    const uint32_t r_end = min(C.n, r_begin + P.rows_per_chunk);
    const uint32_t kp = P.kp;

    // per-(query, chunk) list storage
    auto list_d = [&](uint32_t ql, int which) { return part_d + (((size_t)(q0 + ql) * P.nchunks + chunk) * 2 + which) * kp; };
    auto list_i = [&](uint32_t ql, int which) { return part_i + (((size_t)(q0 + ql) * P.nchunks + chunk) * 2 + which) * kp; };

    for (int ql = warp; ql < BQ; ql += EX_THREADS / 32) {
        if (q0 + ql < Q.nq) for (uint32_t i = lane; i < 2 * kp; i += 32) { list_d(ql, 0)[i] = INFINITY; list_i(ql, 0)[i] = NONE; }
    }
    if (t < BQ) { thr[t] = INFINITY; ccnt[t] = 0; par[t] = 0; }
    __syncthreads();

    for (uint32_t r0 = r_begin; r0 < r_end; r0 += BC) {
        float acc[8][8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

        for (uint32_t kc = 0; kc < Q.dpad; kc += KC) {
            // stage KC dims of 128 queries and 128 rows, transposed to [k][row]
#pragma unroll
            for (int rep = 0; rep < 2; ++rep) {
                const int idx = t + rep * EX_THREADS;       // 0..511
                const int row = idx >> 2, kq = (idx & 3) * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f), w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (q0 + row < Q.nq && kc + kq < Q.dpad) v = *reinterpret_cast<const float4*>(Q.q + (size_t)(q0 + row) * Q.stride + kc + kq);
                if (r0 + row < r_end && kc + kq < Q.dpad) w = __ldg(reinterpret_cast<const float4*>(C.vecs + (size_t)(r0 + row) * C.row_stride + kc + kq));
                qs[kq + 0][row] = v.x; qs[kq + 1][row] = v.y; qs[kq + 2][row] = v.z; qs[kq + 3][row] = v.w;
                cs[kq + 0][row] = w.x; cs[kq + 1][row] = w.y; cs[kq + 2][row] = w.z; cs[kq + 3][row] = w.w;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) {
                const float4 qa = *reinterpret_cast<const float4*>(&qs[kk][ty * 4]);
                const float4 qb = *reinterpret_cast<const float4*>(&qs[kk][64 + ty * 4]);
                const float4 ca = *reinterpret_cast<const float4*>(&cs[kk][tx * 4]);
                const float4 cb = *reinterpret_cast<const float4*>(&cs[kk][64 + tx * 4]);
                const float qv[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
                const float cv[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (METRIC == L2SQ) { const float d = qv[i] - cv[j]; acc[i][j] = fmaf(d, d, acc[i][j]); }
                        else acc[i][j] = fmaf(qv[i], cv[j], acc[i][j]);
                    }
            }
            __syncthreads();
        }

        // finish the metric and find which pairs can enter their query's list
        unsigned long long pend = 0ull;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int ql = (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
            const bool qok = q0 + ql < Q.nq;
            const float qnv = (METRIC == COS && qok) ? qn[q0 + ql] : 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int cl = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
                const uint32_t r = r0 + cl;
                if (METRIC != L2SQ) {
                    const float vn = (METRIC == COS && r < r_end) ? C.sqn[(size_t)r * C.row_step] : 0.f;
                    acc[i][j] = finish_metric(METRIC, acc[i][j], qnv, vn);
                }
                if (qok && r < r_end && !(acc[i][j] > thr[ql])) pend |= 1ull << (i * 8 + j);
            }
        }
        // drop tombstones / self matches (rare path: only pairs that pass the threshold)
        if (pend) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) if (pend >> (i * 8 + j) & 1ull) {
                    const int ql = (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
                    const int cl = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
                    const uint32_t slot = C.first_slot + (r0 + cl) * C.row_step;
                    bool drop = C.deleted && ((C.deleted[slot >> 5] >> (slot & 31u)) & 1u);
                    if (Q.self_step && slot == Q.self_first + (q0 + ql) * Q.self_step) drop = true;
                    if (drop) pend &= ~(1ull << (i * 8 + j));
                }
        }
        // append / merge rounds until every pending pair is placed or rejected
        for (;;) {
            if (pend) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) if (pend >> (i * 8 + j) & 1ull) {
                        const int ql = (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
                        const int cl = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
                        if (acc[i][j] > thr[ql]) { pend &= ~(1ull << (i * 8 + j)); continue; }
                        const uint32_t pos = atomicAdd(&ccnt[ql], 1u);
                        if (pos < CB) {
                            cbuf_d[ql][pos] = acc[i][j]; cbuf_i[ql][pos] = C.first_slot + (r0 + cl) * C.row_step;
                            pend &= ~(1ull << (i * 8 + j));
                        }
                    }
            }
            __syncthreads();
            for (int ql = warp; ql < BQ; ql += EX_THREADS / 32) {
                const uint32_t cnt = min(ccnt[ql], (uint32_t)CB);
                if (cnt == 0) continue;
                SList L; const int w = par[ql];
                L.ld = list_d(ql, w); L.li = list_i(ql, w); L.ld2 = list_d(ql, w ^ 1); L.li2 = list_i(ql, w ^ 1);
                L.cap = (int)k; L.capp = (int)kp;
                const float d = lane < (int)cnt ? cbuf_d[ql][lane] : INFINITY;
                const uint32_t id = lane < (int)cnt ? cbuf_i[ql][lane] : NONE;
                __syncwarp();
                slist_merge(L, d, id, scr_id[warp], scr_d[warp], less);
                __syncwarp();
                if (lane == 0) { par[ql] = (uint8_t)(w ^ 1); thr[ql] = L.ld[k - 1]; ccnt[ql] = 0; }
            }
            __threadfence_block();
            if (!__syncthreads_or(pend != 0ull)) break;
        }
    }
    if (t < BQ && q0 + t < Q.nq) par_out[(size_t)(q0 + t) * P.nchunks + chunk] = par[t];
}

// ------------------------------------------------------------------ K1s: streaming exact scan --
// Few queries (nq <= 16) against a large corpus is HBM-bound (4 N d bytes per pass,
// 2 nq N d flop): the 128 x 128 tile kernel above wastes 127/128 of its tile at nq = 1 and the tensor
// path pads to a 128-query tile. Here one thread owns one corpus row and up to 8 queries ride along:
// a block of 256 threads streams tiles of 256 rows x 32 floats through a ring of TMA stages
// (cp.async.bulk.tensor.2d, 128 B swizzle so that thread-per-row float4 reads are bank-conflict free),
// every (query, row) accumulator is one sequential fmaf chain over the dimensions in index order
// (packed f32x2: the two halves belong to two QUERIES, never to two dimensions), so ids and distance
// bits are the oracle's. Per-query top-k lists of the block live in shared memory (threshold + append
// + warp merge, as in k_exact_tiles) and are written once at the end in the per-chunk layout
// k_exact_final folds.
constexpr int SR = 256, SKC = 32, ST_THREADS = 256, ST_MAX_STAGES = 6, SCB = 32;
constexpr size_t ST_STAGE_BYTES = (size_t)SR * SKC * 4;

__device__ __forceinline__ void st_tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

// dynamic shared memory: [<= 1023 B alignment slack] stages | q pairs [NP][dq] float2 | lists [2 NP][2][kp] (d, id) | ctrl
__host__ __device__ inline size_t stream_smem_bytes(uint32_t np, uint32_t dpad, uint32_t kp, uint32_t nst) {
    const size_t dq = (dpad + SKC - 1) / SKC * SKC;
    return 1024 + nst * ST_STAGE_BYTES + (size_t)np * dq * 8 + (size_t)2 * np * 2 * kp * 8 + (size_t)2 * np * SCB * 8 + 3072;
}

template <int METRIC, int NP, class Less>
__global__ void __launch_bounds__(ST_THREADS, 1)
k_exact_stream(const __grid_constant__ CUtensorMap cmap, ExactCorpus C, ExactQueries Q, uint32_t q0, uint32_t nqv, ExactPlan P, uint32_t k, uint32_t nst,
               const float* __restrict__ qn, float* part_d, uint32_t* part_i, uint8_t* par_out, Less less) {
    constexpr int NQ = 2 * NP;
    extern __shared__ __align__(16) unsigned char st_raw[];
    unsigned char* sm = st_raw + ((1024u - (smem_u32(st_raw) & 1023u)) & 1023u);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const uint32_t kp = P.kp, dq = (Q.dpad + SKC - 1) / SKC * SKC, nkc = dq / SKC;
    unsigned char* p = sm;
    const uint32_t stage_s = smem_u32(p); const unsigned char* stage_g = p; p += nst * ST_STAGE_BYTES;
    float2* qs = reinterpret_cast<float2*>(p); p += (size_t)NP * dq * 8;
    float* lst_d = reinterpret_cast<float*>(p); p += (size_t)NQ * 2 * kp * 4;
    uint32_t* lst_i = reinterpret_cast<uint32_t*>(p); p += (size_t)NQ * 2 * kp * 4;
    float* cbuf_d = reinterpret_cast<float*>(p); p += NQ * SCB * 4;
    uint32_t* cbuf_i = reinterpret_cast<uint32_t*>(p); p += NQ * SCB * 4;
    float* thr = reinterpret_cast<float*>(p); p += NQ * 4;
    uint32_t* ccnt = reinterpret_cast<uint32_t*>(p); p += NQ * 4;
    uint32_t* par = reinterpret_cast<uint32_t*>(p); p += NQ * 4;
    float* qnv = reinterpret_cast<float*>(p); p += NQ * 4;
    p = sm + ((((size_t)(p - sm)) + 7) & ~(size_t)7);
    const uint32_t bar0 = smem_u32(p); p += ST_MAX_STAGES * 8;
    uint32_t* scr_id = reinterpret_cast<uint32_t*>(p); p += (ST_THREADS / 32) * 32 * 4;
    float* scr_d = reinterpret_cast<float*>(p);

    const uint32_t chunk = blockIdx.x;
    const uint32_t r_begin = chunk * P.rows_per_chunk, r_end = min(C.n, r_begin + P.rows_per_chunk);
    const uint32_t ntiles = r_end > r_begin ? (r_end - r_begin + SR - 1) / SR : 0, total = ntiles * nkc;

    if (t < (int)nst) mbar_init(bar0 + 8u * t, 1);
    // queries of this pass, interleaved in pairs: qs[pr][c] = (q[2 pr][c], q[2 pr + 1][c]); absent queries and c >= dpad are zero
    for (uint32_t i = t; i < (uint32_t)NP * dq; i += ST_THREADS) {
        const uint32_t pr = i / dq, c = i - pr * dq;
        const uint32_t qa = 2 * pr, qb = 2 * pr + 1;
        float2 v = make_float2(0.f, 0.f);
        if (c < Q.dpad) {
            if (qa < nqv) v.x = Q.q[(size_t)(q0 + qa) * Q.stride + c];
            if (qb < nqv) v.y = Q.q[(size_t)(q0 + qb) * Q.stride + c];
        }
        qs[i] = v;
    }
    for (uint32_t i = t; i < (uint32_t)NQ * 2 * kp; i += ST_THREADS) { lst_d[i] = INFINITY; lst_i[i] = NONE; }
    if (t < NQ) { thr[t] = INFINITY; ccnt[t] = 0; par[t] = 0; qnv[t] = (METRIC == COS && (uint32_t)t < nqv) ? qn[q0 + t] : 0.f; }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    auto issue = [&](uint32_t it, uint32_t slot) {   // thread 0 only
        const uint32_t tile = it / nkc, kc = it - tile * nkc, bar = bar0 + 8u * slot;
        mbar_arrive_expect_tx(bar, (uint32_t)ST_STAGE_BYTES);
        st_tma_load_2d(stage_s + slot * (uint32_t)ST_STAGE_BYTES, &cmap, bar, (int)(kc * SKC), (int)(r_begin + tile * SR));
    };
    if (t == 0) for (uint32_t it = 0; it + 1 < nst && it < total; ++it) issue(it, it);

    auto try_append = [&](const float (&dist)[NQ], unsigned& pend, uint32_t slot_id) {
        if (!pend) return;
#pragma unroll
        for (int b = 0; b < NQ; ++b) if (pend >> b & 1u) {
            if (dist[b] > thr[b]) { pend &= ~(1u << b); continue; }
            const uint32_t pos = atomicAdd(&ccnt[b], 1u);
            if (pos < SCB) { cbuf_d[b * SCB + pos] = dist[b]; cbuf_i[b * SCB + pos] = slot_id; pend &= ~(1u << b); }
        }
    };
    auto merge_all = [&]() {   // warp b folds query b's buffer into its list; callers bracket it with block barriers
        for (int b = warp; b < NQ; b += ST_THREADS / 32) {
            const uint32_t cnt = min(ccnt[b], (uint32_t)SCB);
            if (cnt == 0) continue;
            SList L; const uint32_t w = par[b];
            L.ld = lst_d + ((size_t)b * 2 + w) * kp; L.li = lst_i + ((size_t)b * 2 + w) * kp;
            L.ld2 = lst_d + ((size_t)b * 2 + (w ^ 1u)) * kp; L.li2 = lst_i + ((size_t)b * 2 + (w ^ 1u)) * kp;
            L.cap = (int)k; L.capp = (int)kp;
            const float d = lane < (int)cnt ? cbuf_d[b * SCB + lane] : INFINITY;
            const uint32_t id = lane < (int)cnt ? cbuf_i[b * SCB + lane] : NONE;
            __syncwarp();
            slist_merge(L, d, id, scr_id + warp * 32, scr_d + warp * 32, less);
            __syncwarp();
            if (lane == 0) { par[b] = w ^ 1u; thr[b] = L.ld[k - 1]; ccnt[b] = 0; }
        }
    };
    float2 acc[NP];
#pragma unroll
    for (int b = 0; b < NP; ++b) acc[b] = make_float2(0.f, 0.f);
    const float2 m1 = make_float2(-1.f, -1.f);
    uint32_t tile = 0, kc = 0, slot = 0, phase = 0;
    for (uint32_t it = 0; it < total; ++it) {
        // slot (it - 1) % nst was released by the barrier that ended iteration it - 1
        if (t == 0 && it + nst - 1 < total) issue(it + nst - 1, slot == 0 ? nst - 1 : slot - 1);
        mbar_wait(bar0 + 8u * slot, phase);
        const float4* xs = reinterpret_cast<const float4*>(stage_g + slot * ST_STAGE_BYTES + (size_t)t * (SKC * 4));
        const float4* q4 = reinterpret_cast<const float4*>(qs + kc * SKC);   // pair pr at q4 + pr * dq / 2
#pragma unroll
        for (int j = 0; j < SKC / 4; ++j) {
            const float4 x = xs[j ^ (t & 7)];
#pragma unroll
            for (int b = 0; b < NP; ++b) {
                const float4 qa = q4[(size_t)b * (dq / 2) + 2 * j], qb = q4[(size_t)b * (dq / 2) + 2 * j + 1];
                if (METRIC == L2SQ) {
                    float2 d0 = __ffma2_rn(make_float2(x.x, x.x), m1, make_float2(qa.x, qa.y));
                    float2 d1 = __ffma2_rn(make_float2(x.y, x.y), m1, make_float2(qa.z, qa.w));
                    float2 d2 = __ffma2_rn(make_float2(x.z, x.z), m1, make_float2(qb.x, qb.y));
                    float2 d3 = __ffma2_rn(make_float2(x.w, x.w), m1, make_float2(qb.z, qb.w));
                    acc[b] = __ffma2_rn(d0, d0, acc[b]); acc[b] = __ffma2_rn(d1, d1, acc[b]);
                    acc[b] = __ffma2_rn(d2, d2, acc[b]); acc[b] = __ffma2_rn(d3, d3, acc[b]);
                } else {
                    acc[b] = __ffma2_rn(make_float2(qa.x, qa.y), make_float2(x.x, x.x), acc[b]);
                    acc[b] = __ffma2_rn(make_float2(qa.z, qa.w), make_float2(x.y, x.y), acc[b]);
                    acc[b] = __ffma2_rn(make_float2(qb.x, qb.y), make_float2(x.z, x.z), acc[b]);
                    acc[b] = __ffma2_rn(make_float2(qb.z, qb.w), make_float2(x.w, x.w), acc[b]);
                }
            }
        }
        if (++kc == nkc) {   // the tile's rows are complete: finish the metric, offer them to the lists
            kc = 0;
            const uint32_t r = r_begin + tile * SR + t;
            const uint32_t slot_id = C.first_slot + r * C.row_step;
            ++tile;
            float dist[NQ];
            unsigned pend = 0u;
            const bool rok = r < r_end;
            const float vn = (METRIC == COS && rok) ? C.sqn[(size_t)r * C.row_step] : 0.f;
#pragma unroll
            for (int b = 0; b < NQ; ++b) {
                const float a = (b & 1) ? acc[b >> 1].y : acc[b >> 1].x;
                dist[b] = finish_metric(METRIC, a, qnv[b], vn);
                if (rok && (uint32_t)b < nqv && !(dist[b] > thr[b])) pend |= 1u << b;
            }
#pragma unroll
            for (int b = 0; b < NP; ++b) acc[b] = make_float2(0.f, 0.f);
            if (pend) {   // rare: tombstones / the query's own row
                bool drop = C.deleted && ((C.deleted[slot_id >> 5] >> (slot_id & 31u)) & 1u);
                if (drop) pend = 0u;
                if (Q.self_step) {
#pragma unroll
                    for (int b = 0; b < NQ; ++b) if (slot_id == Q.self_first + (q0 + b) * Q.self_step) pend &= ~(1u << b);
                }
            }
            // Lazy lists: candidates are only appended to the per-query buffers; the buffers are folded into the sorted
            // lists (and the thresholds refreshed) when one overflows, so a block pays a merge round once per ~32
            // accepted rows of a query instead of once per tile with any accepted row.
            try_append(dist, pend, slot_id);
            while (__syncthreads_or(pend != 0u)) {
                merge_all();
                __syncthreads();
                try_append(dist, pend, slot_id);
            }
        }
        __syncthreads();   // every thread is done with this stage: thread 0 may refill it
        if (++slot == nst) { slot = 0; phase ^= 1u; }
    }
    merge_all();   // what is still waiting in the buffers (the loop's last barrier made it visible)
    __syncthreads();
    // the block's lists -> the per-(query, chunk) layout of k_exact_final (parity 0)
    for (uint32_t b = 0; b < nqv; ++b) {
        const size_t lb = (((size_t)(q0 + b) * P.nchunks + chunk) * 2) * kp;
        const uint32_t w = par[b];
        for (uint32_t i = t; i < kp; i += ST_THREADS) { part_d[lb + i] = lst_d[((size_t)b * 2 + w) * kp + i]; part_i[lb + i] = lst_i[((size_t)b * 2 + w) * kp + i]; }
        if (t == 0) par_out[(size_t)(q0 + b) * P.nchunks + chunk] = 0;
    }
}

// The streaming scan applies to contiguous corpora (the search-side call) with few queries.
static bool stream_applies(const ExactCorpus& c, const ExactQueries& q, uint32_t k, const DeviceInfo& dev) {
    const char* env = getenv("VSANN_EXACT_STREAM");   // 0 = off, 1 = whenever it fits, default: by size
    const int mode = env ? atoi(env) : -1;
    if (mode == 0) return false;
    if (c.row_step != 1 || c.row_stride != q.dpad || q.nq == 0 || q.nq > 16 || k == 0 || k > 256) return false;
    if ((q.dpad & 3u) || ((uintptr_t)c.vecs & 15u)) return false;
    const uint32_t kp = (k + 31u) & ~31u;
    if (stream_smem_bytes(1, q.dpad, kp, 3) > dev.smem_per_block) return false;
    return mode == 1 || (uint64_t)c.n * q.dpad >= (1ull << 22);
}
bool exact_stream_supported(const ExactCorpus& c, const ExactQueries& q, uint32_t k, const DeviceInfo& dev) { return stream_applies(c, q, k, dev); }

template <int METRIC, int NP, class Less>
static cudaError_t launch_stream_pass(const CUtensorMap& cmap, const ExactCorpus& c, const ExactQueries& q, uint32_t q0, uint32_t nqv, const ExactPlan& P,
                                      uint32_t k, uint32_t nst, const float* qn, float* part_d, uint32_t* part_i, uint8_t* par, size_t smem, cudaStream_t st, Less less) {
    cudaError_t e = cudaFuncSetAttribute(k_exact_stream<METRIC, NP, Less>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_exact_stream<METRIC, NP, Less><<<P.nchunks, ST_THREADS, smem, st>>>(cmap, c, q, q0, nqv, P, k, nst, qn, part_d, part_i, par, less);
    return cudaGetLastError();
}

template <int METRIC, class Less>
static cudaError_t run_stream_metric(const CUtensorMap& cmap, const ExactCorpus& c, const ExactQueries& q, const ExactPlan& P, uint32_t k, const float* qn,
                                     float* part_d, uint32_t* part_i, uint8_t* par, const DeviceInfo& dev, cudaStream_t st, Less less) {
    // queries per pass: as many pairs as shared memory holds next to the stages (1, 2 or 4 pairs)
    uint32_t np_max = 4;
    while (np_max > 1 && stream_smem_bytes(np_max, q.dpad, P.kp, 3) > dev.smem_per_block) np_max >>= 1;
    for (uint32_t q0 = 0; q0 < q.nq;) {
        const uint32_t left = q.nq - q0;
        uint32_t np = 1; while (np < np_max && 2 * np < left) np <<= 1;
        const uint32_t nqv = std::min<uint32_t>(left, 2 * np);
        uint32_t nst = ST_MAX_STAGES;   // as deep a ring as shared memory holds: the scan is latency x bandwidth bound
        while (nst > 3 && stream_smem_bytes(np, q.dpad, P.kp, nst) > dev.smem_per_block) --nst;
        if (const char* e = getenv("VSANN_STREAM_STAGES")) nst = std::max<uint32_t>(2u, std::min<uint32_t>(nst, (uint32_t)atoi(e)));   // A/B switch
        const size_t smem = stream_smem_bytes(np, q.dpad, P.kp, nst);
        cudaError_t e;
        if (np == 1) e = launch_stream_pass<METRIC, 1, Less>(cmap, c, q, q0, nqv, P, k, nst, qn, part_d, part_i, par, smem, st, less);
        else if (np == 2) e = launch_stream_pass<METRIC, 2, Less>(cmap, c, q, q0, nqv, P, k, nst, qn, part_d, part_i, par, smem, st, less);
        else e = launch_stream_pass<METRIC, 4, Less>(cmap, c, q, q0, nqv, P, k, nst, qn, part_d, part_i, par, smem, st, less);
        if (e != cudaSuccess) return e;
        q0 += nqv;
    }
    return cudaSuccess;
}

// One warp per (query, group of chunks): fold the per-chunk lists into one ascending top-k.
// groups == 1: the final answer goes to out_*; groups > 1 (first level of a two-level fold, streaming scan only,
// whose lists all have parity 0): group g of query q leaves its list in the parity-1 buffer of chunk g.
// in_par >= 0: every input list has that parity (second level), else par[] says which.
template <class Less>
__global__ void k_exact_final(ExactPlan P, uint32_t nq, uint32_t k, uint32_t out_ld, float* part_d, uint32_t* part_i, const uint8_t* par,
                              uint32_t c_count, uint32_t groups, int in_par,
                              const uint64_t* __restrict__ slot_key, uint32_t* out_slots, uint64_t* out_keys, float* out_dist, Less less) {
    extern __shared__ __align__(16) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    const uint32_t w = blockIdx.x * wpb + warp;
    if (w >= nq * groups) return;
    const uint32_t q = w / groups, grp = w - q * groups;
    const uint32_t per = (c_count + groups - 1) / groups, c_begin = grp * per, c_end = min(c_count, c_begin + per);
    const uint32_t kp = P.kp;
    unsigned char* base = sm + (size_t)warp * (kp * 16 + 256);
    SList L;
    L.ld = (float*)base; L.ld2 = L.ld + kp; L.li = (uint32_t*)(L.ld2 + kp); L.li2 = L.li + kp;
    uint32_t* scr_id = L.li2 + kp; float* scr_d = (float*)(scr_id + 32);
    L.cap = (int)k; L.capp = (int)kp;
    slist_init(L);
    for (uint32_t c = c_begin; c < c_end; ++c) {
        const uint32_t pw = in_par >= 0 ? (uint32_t)in_par : par[(size_t)q * P.nchunks + c];
        const size_t lb = (((size_t)q * P.nchunks + c) * 2 + pw) * kp;
        for (uint32_t b = 0; b < kp; b += 32) {
            const float d = part_d[lb + b + lane]; const uint32_t id = part_i[lb + b + lane];
            if (__ballot_sync(0xffffffffu, id != NONE) == 0u) break;
            // chunk lists ascend: once the group's smallest entry cannot enter the full list, neither can the rest
            const float d0 = __shfl_sync(0xffffffffu, d, 0); const uint32_t i0 = __shfl_sync(0xffffffffu, id, 0);
            const float wd = L.ld[k - 1]; const uint32_t wi = L.li[k - 1];
            if (wi != NONE && !less(d0, i0, wd, wi)) break;
            slist_merge(L, d, id, scr_id, scr_d, less);
        }
    }
    if (groups > 1) {
        const size_t ob = (((size_t)q * P.nchunks + grp) * 2 + 1) * kp;
        for (uint32_t i = lane; i < kp; i += 32) { part_d[ob + i] = i < k ? L.ld[i] : INFINITY; part_i[ob + i] = i < k ? L.li[i] : NONE; }
        return;
    }
    for (uint32_t i = lane; i < out_ld; i += 32) {
        const uint32_t id = i < k ? L.li[i] : NONE;
        if (out_slots) out_slots[(size_t)q * out_ld + i] = id;
        if (out_keys) out_keys[(size_t)q * out_ld + i] = id == NONE ? 0xFFFFFFFFFFFFFFFFull : slot_key[id];
        if (out_dist) out_dist[(size_t)q * out_ld + i] = i < k ? L.ld[i] : INFINITY;
    }
}

template <class Less>
static cudaError_t run_exact(const ExactCorpus& c, const ExactQueries& q, int metric, uint32_t k, uint32_t out_ld, uint32_t* out_slots, uint64_t* out_keys,
                             float* out_dist, const ExactScratch& s, const DeviceInfo& dev, cudaStream_t st, Less less) {
    ExactPlan P = plan_exact(q.nq, c.n, k, dev.sm_count);
    const size_t mc = max_chunks(q.nq, dev.sm_count);
    if (P.nchunks > mc || s.bytes < exact_scratch_bytes(q.nq, k, dev.sm_count)) return cudaErrorInvalidValue;
    unsigned char* p = (unsigned char*)s.ptr;
    float* qn = (float*)p; p += ((size_t)q.nq * 4 + 255) & ~(size_t)255;
    uint8_t* par = (uint8_t*)p; p += ((size_t)q.nq * mc + 255) & ~(size_t)255;
    float* part_d = (float*)p; p += (size_t)q.nq * mc * 2 * P.kp * 4;
    p = (unsigned char*)(((uintptr_t)p + 255) & ~(uintptr_t)255);
    uint32_t* part_i = (uint32_t*)p;
    if (metric == COS) k_query_norms<<<(q.nq + 127) / 128, 128, 0, st>>>(q, qn);
    cudaError_t e = cudaSuccess;
    CUtensorMap cmap;
    bool streamed = false;
    if (stream_applies(c, q, k, dev) && make_tmap_f32(&cmap, c.vecs, c.n, q.dpad, c.row_stride, SKC, SR, CU_TENSOR_MAP_SWIZZLE_128B)) {
        // K1s: HBM-bound streaming scan for a handful of queries; one block per SM, one chunk per block
        streamed = true;
        const uint32_t tiles = (c.n + SR - 1) / SR, want = std::min<uint32_t>(tiles, (uint32_t)dev.sm_count);
        P.rows_per_chunk = ((tiles + want - 1) / want) * SR;
        P.nchunks = (c.n + P.rows_per_chunk - 1) / P.rows_per_chunk;
        switch (metric) {
            case L2SQ: e = run_stream_metric<L2SQ, Less>(cmap, c, q, P, k, qn, part_d, part_i, par, dev, st, less); break;
            case IP: e = run_stream_metric<IP, Less>(cmap, c, q, P, k, qn, part_d, part_i, par, dev, st, less); break;
            default: e = run_stream_metric<COS, Less>(cmap, c, q, P, k, qn, part_d, part_i, par, dev, st, less); break;
        }
    } else {
        const dim3 grid(P.nchunks, P.nqt);
        switch (metric) {
            case L2SQ: k_exact_tiles<L2SQ, Less><<<grid, EX_THREADS, 0, st>>>(c, q, P, k, qn, part_d, part_i, par, less); break;
            case IP: k_exact_tiles<IP, Less><<<grid, EX_THREADS, 0, st>>>(c, q, P, k, qn, part_d, part_i, par, less); break;
            default: k_exact_tiles<COS, Less><<<grid, EX_THREADS, 0, st>>>(c, q, P, k, qn, part_d, part_i, par, less); break;
        }
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) return e;
    const uint32_t wpb = 4; const size_t smem = (size_t)wpb * (P.kp * 16 + 256);
    if (smem > 48 * 1024) {
        e = cudaFuncSetAttribute(k_exact_final<Less>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    // few queries after the streaming scan (one list per SM, all parity 0): fold in two levels so that the
    // chain of dependent merges is ~2 sqrt(chunks) long instead of `chunks`
    uint32_t groups = 1;
    if (streamed && P.nchunks >= 32) { groups = 1; while (groups * groups < P.nchunks) ++groups; }
    if (groups > 1) {
        k_exact_final<Less><<<(q.nq * groups + wpb - 1) / wpb, wpb * 32, smem, st>>>(P, q.nq, k, out_ld, part_d, part_i, par, P.nchunks, groups, -1,
                                                                                   c.slot_key, nullptr, nullptr, nullptr, less);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        k_exact_final<Less><<<(q.nq + wpb - 1) / wpb, wpb * 32, smem, st>>>(P, q.nq, k, out_ld, part_d, part_i, par, groups, 1, 1,
                                                                          c.slot_key, out_slots, out_keys, out_dist, less);
        return cudaGetLastError();
    }
    k_exact_final<Less><<<(q.nq + wpb - 1) / wpb, wpb * 32, smem, st>>>(P, q.nq, k, out_ld, part_d, part_i, par, P.nchunks, 1, -1,
                                                                      c.slot_key, out_slots, out_keys, out_dist, less);
    return cudaGetLastError();
}

__global__ void k_fill_empty(size_t n, uint32_t* out_slots, uint64_t* out_keys, float* out_dist) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if (out_slots) out_slots[i] = NONE;
        if (out_keys) out_keys[i] = 0xFFFFFFFFFFFFFFFFull;
        if (out_dist) out_dist[i] = INFINITY;
    }
}

cudaError_t launch_exact(const ExactCorpus& c, const ExactQueries& q, int metric, uint32_t k, uint32_t out_ld,
                         uint32_t* out_slots, uint64_t* out_keys, float* out_dist,
                         const ExactScratch& scratch, const DeviceInfo& dev, cudaStream_t st) {
    if (q.nq == 0 || k == 0) return cudaSuccess;
    if (out_ld < k) return cudaErrorInvalidValue;
    if (c.n == 0) {
        k_fill_empty<<<dev.sm_count, 256, 0, st>>>((size_t)q.nq * out_ld, out_slots, out_keys, out_dist);
        return cudaGetLastError();
    }
    if (c.slot_key) return run_exact(c, q, metric, k, out_ld, out_slots, out_keys, out_dist, scratch, dev, st, LessByKey{c.slot_key, c.n_slots});
    return run_exact(c, q, metric, k, out_ld, out_slots, out_keys, out_dist, scratch, dev, st, LessById());
}

}  // namespace vsann
