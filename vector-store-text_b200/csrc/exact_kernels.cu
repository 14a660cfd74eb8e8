// exact_kernels.cu — K1/K2: exact (brute-force) k-NN, FP32, bit-identical to the oracle.
//
// The north star's "reference's exact search" has no call site in the reference;
// it is defined by oracle/oracle_vsann.cpp exact_search: f32 distance with a sequential fmaf
// chain over the dimensions in index order, results ordered by (distance asc, key asc).
// This kernel keeps that exact accumulation order — every (query, row) pair owns one
// accumulator that is updated for k = 0, 1, ..., d-1 in order — so ids AND distances match the
// oracle bit for bit. It is the register-tiled FP32 form (128 x 128 tile, 8 x 8 per thread);
// the same tiles also produce the intra-batch neighbours of an insertion batch (K4).
#include "common.cuh"
#include "kernels.h"

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

// One warp per query: fold the per-chunk lists into the final ascending top-k.
template <class Less>
__global__ void k_exact_final(ExactPlan P, uint32_t nq, uint32_t k, uint32_t out_ld, const float* part_d, const uint32_t* part_i, const uint8_t* par,
                              const uint64_t* __restrict__ slot_key, uint32_t* out_slots, uint64_t* out_keys, float* out_dist, Less less) {
    extern __shared__ __align__(16) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    const uint32_t q = blockIdx.x * wpb + warp;
    if (q >= nq) return;
    const uint32_t kp = P.kp;
    unsigned char* base = sm + (size_t)warp * (kp * 16 + 256);
    SList L;
    L.ld = (float*)base; L.ld2 = L.ld + kp; L.li = (uint32_t*)(L.ld2 + kp); L.li2 = L.li + kp;
    uint32_t* scr_id = L.li2 + kp; float* scr_d = (float*)(scr_id + 32);
    L.cap = (int)k; L.capp = (int)kp;
    slist_init(L);
    for (uint32_t c = 0; c < P.nchunks; ++c) {
        const size_t lb = (((size_t)q * P.nchunks + c) * 2 + par[(size_t)q * P.nchunks + c]) * kp;
        for (uint32_t b = 0; b < kp; b += 32) {
            const float d = part_d[lb + b + lane]; const uint32_t id = part_i[lb + b + lane];
            if (__ballot_sync(0xffffffffu, id != NONE) == 0u) break;
            slist_merge(L, d, id, scr_id, scr_d, less);
        }
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
    const ExactPlan P = plan_exact(q.nq, c.n, k, dev.sm_count);
    const size_t mc = max_chunks(q.nq, dev.sm_count);
    if (P.nchunks > mc || s.bytes < exact_scratch_bytes(q.nq, k, dev.sm_count)) return cudaErrorInvalidValue;
    unsigned char* p = (unsigned char*)s.ptr;
    float* qn = (float*)p; p += ((size_t)q.nq * 4 + 255) & ~(size_t)255;
    uint8_t* par = (uint8_t*)p; p += ((size_t)q.nq * mc + 255) & ~(size_t)255;
    float* part_d = (float*)p; p += (size_t)q.nq * mc * 2 * P.kp * 4;
    p = (unsigned char*)(((uintptr_t)p + 255) & ~(uintptr_t)255);
    uint32_t* part_i = (uint32_t*)p;
    if (metric == COS) k_query_norms<<<(q.nq + 127) / 128, 128, 0, st>>>(q, qn);
    const dim3 grid(P.nchunks, P.nqt);
    switch (metric) {
        case L2SQ: k_exact_tiles<L2SQ, Less><<<grid, EX_THREADS, 0, st>>>(c, q, P, k, qn, part_d, part_i, par, less); break;
        case IP: k_exact_tiles<IP, Less><<<grid, EX_THREADS, 0, st>>>(c, q, P, k, qn, part_d, part_i, par, less); break;
        default: k_exact_tiles<COS, Less><<<grid, EX_THREADS, 0, st>>>(c, q, P, k, qn, part_d, part_i, par, less); break;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const uint32_t wpb = 4; const size_t smem = (size_t)wpb * (P.kp * 16 + 256);
    if (smem > 48 * 1024) {
        e = cudaFuncSetAttribute(k_exact_final<Less>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_exact_final<Less><<<(q.nq + wpb - 1) / wpb, wpb * 32, smem, st>>>(P, q.nq, k, out_ld, part_d, part_i, par, c.slot_key, out_slots, out_keys, out_dist, less);
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
