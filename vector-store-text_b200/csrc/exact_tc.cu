// exact_tc.cu — K1 on the 5th-generation tensor cores: brute-force k-NN as a TF32 tcgen05 GEMM with
// TMA-fed tiles and TMEM accumulators, followed by an exact FP32 re-rank (K2) so that the ids AND
// the distance bits are those of the oracle's exact search (oracle/oracle_vsann.cpp exact_search;
// "the reference's exact search", SURVEY.md §8c — usearch::Index::search at usearch.rs:276 run
// exhaustively).
//
//   1. k_tc_scan<SAMPLE>  D = Q · X^T on every 32nd corpus tile; every query keeps its R smallest keys
//   2. k_tc_threshold     tau[q] = r-th smallest sampled key  (expected |{x : key <= tau}| = 32 r >> k)
//   3. k_tc_scan<EMIT>    full scan; (q, slot) pairs with key <= tau[q] are appended to cand[q][]
//   4. k_tc_rerank        exact sequential-fmaf distance of every candidate, top-k by (distance, key);
//                         proves the answer: every row that was NOT emitted has approximate distance
//                         > tau, hence exact distance > tau - eps >= the k-th exact distance found
//   5. queries whose proof fails (ties wider than eps, candidate overflow, < k live candidates) are
//      answered by the FP32 tile kernel (exact_kernels.cu). There is no CPU path.
//
// key = a_j * dot + b_j with per-row (a_j, b_j): L2SQ (-2, |x|^2), IP (-1, 0), COS (-1/|x|, 0);
// tombstoned rows and rows past the end get (0, +inf). dist = f(key) is monotone per query.
//
// CTA = 128 queries (UMMA M) x one corpus slice, tiles of 256 rows (UMMA N), K in blocks of 32 floats
// (one 128 B swizzle row). Warps 0..E-1: epilogue (TMEM lane quarter = warp % 4, column part = warp / 4; E = 8 sampling, 16 full scan),
// warp E: TMA producer, warp E+1: MMA issuer + TMEM allocator, warp E+2: per-row coefficients. 4-stage smem ring, 2 TMEM accumulators of 256 columns.
// Queries of up to 128 dimensions (4 K blocks) stay RESIDENT in shared memory for the CTA's lifetime: the ring then carries corpus
// tiles only and the L2 -> SM traffic per tile drops from 192 KB to 128 KB (the scan ran at the L2's delivery rate: 62 GB per
// 10k x 1M x 128 scan = 12.1 TB/s, profiles/r02m). The coefficient warp runs up to 8 tiles ahead of the epilogue through its own
// ring (it waited for the epilogue to free one of two buffers and then for its global loads: 15 % of the epilogue's samples), and
// the epilogue never waits for an atomic: the ids of a chunk are written when the next chunk with candidates comes by.
#include <cuda.h>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"
#include "tmap.h"

namespace vsann {
namespace tc {

constexpr int BM = 128, BN = 256, BK = 32, STAGES = 4, UMMA_K = 8;
constexpr int A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4, STAGE_BYTES = A_BYTES + B_BYTES;
// epilogue warps: 8 in the sampling pass (a thread keeps R_MAX keys in registers), 16 in the full scan - its epilogue is a chain of
// TMEM loads, shared-memory loads and dependent minima, bound by latency, not issue slots (36 % busy with 8 warps: profiles/r02m)
constexpr int EPI_WARPS = 8, EPI_WARPS_EMIT = 16;
__host__ __device__ constexpr int epi_warps(bool emit) { return emit ? EPI_WARPS_EMIT : EPI_WARPS; }
__host__ __device__ constexpr int threads_of(bool emit) { return epi_warps(emit) * 32 + 96; }   // + TMA, MMA, coefficient warps
constexpr int R_MAX = 16;              // sampled keys kept per (query, slice)
constexpr int SAMPLE_STRIDE = 32;      // large corpora: every 32nd tile is sampled (small ones: Plan::stride)
constexpr uint32_t TMEM_COLS = 512;
constexpr int CB = 8, CG = 4;           // coefficient ring: buffers, tiles whose loads one round of the coefficient warp has in flight
constexpr int RES_KB = 4;              // query K blocks that stay resident (d <= 128)
constexpr uint32_t R2_MAX = 1024;      // queries the second (proven-threshold) scan serves at most; more than that go to the FP32 fallback

// dynamic shared memory (after rounding the base up to 1024 B)
struct __align__(16) Ctrl {
    unsigned long long full[STAGES], empty[STAGES], tfull[2], tempty[2], cfull[CB], cempty[CB], afull;
    uint32_t tmem_base;
    uint32_t pad[1];
};
constexpr size_t COLV_BYTES = (size_t)CB * BN * sizeof(float2);
static_assert((size_t)RES_KB * A_BYTES + (size_t)STAGES * B_BYTES <= (size_t)STAGES * STAGE_BYTES, "resident layout fits the streaming one");
constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + COLV_BYTES + sizeof(Ctrl);

// ---- PTX wrappers --------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// half a corpus tile, delivered to BOTH CTAs of a 2-CTA cluster (same shared-memory offset, same mbarrier offset in each)
__device__ __forceinline__ void tma_load_2d_mc2(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem]^T, TF32 inputs, FP32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// mbarrier arrive once every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// the same arrive on the mbarrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_mc2(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3) : "memory");
}
// issue a 32-column load of this thread's accumulator row; tmem_wait() before the registers are read
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, float (&v)[32]) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile, 128 B swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO), LBO unused (= 1).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::tf32, FP32 accumulate, A and B K-major, M = 128, N = BN
constexpr uint32_t IDESC_TF32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);


// |q|^2 by the sequential fmaf chain of the oracle (same bits as k_query_norms / k_row_norms)
__global__ void k_query_norms_tc(const float* __restrict__ q, size_t stride, uint32_t dpad, uint32_t nq, float* __restrict__ qn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float* v = q + (size_t)i * stride;
    float acc = 0.f;
    for (uint32_t c = 0; c < dpad; ++c) acc = fmaf(v[c], v[c], acc);
    qn[i] = acc;
}

struct ScanArgs {
    const float* sqn; const uint32_t* deleted;   // sqn by row, tombstones by slot (row r is slot first_slot + r)
    uint32_t first_slot, n, nq, nq_pad, kblocks;
    int metric;
    uint32_t n_tiles, tiles_per_slice, sample_stride;
    float* samp;                                 // SAMPLE: [slices * EPI_WARPS / 4][nq_pad][R_MAX] ascending
    const float* tau; uint32_t* cand; uint32_t* cnt; uint32_t cap;   // EMIT
    const uint32_t* n_active; uint32_t active_cap;                    // second scan: the query count lives on the device (null = nq)
    uint32_t flags;                                                   // A/B switches (VSANN_TC_FLAGS): 1 = queries not resident, 2 = coefficient loads one tile at a time, 4 = ids written at once,
                                                                      // 8 = no epilogue arithmetic, 16 = no TMA loads (8 / 24: rate measurements, results are garbage)
};

// (a, b) of one corpus row: key = a * dot + b. In two halves, so that the coefficient warp can issue the loads of several tiles
// back to back and look at the values afterwards: written as one function with early returns, every row cost two DEPENDENT
// round trips to an L2 that is busy delivering 12 TB/s of tiles (16 in a row per tile, 2.4 us: that, not the tensor pipe or the
// epilogue's arithmetic, set the pace of the scan - profiles/r02m).
struct RowRaw { uint32_t bits; float xn; };
__device__ __forceinline__ RowRaw row_load(const ScanArgs& a, uint32_t r) {
    const uint32_t rr = min(r, a.n - 1u), slot = a.first_slot + rr;
    const uint32_t* del = a.deleted ? a.deleted : reinterpret_cast<const uint32_t*>(a.sqn);     // no tombstones: any readable word
    RowRaw x;
    x.bits = __ldg(del + (a.deleted ? (slot >> 5) : 0u));
    x.xn = __ldg(a.sqn + rr);
    return x;
}
__device__ __forceinline__ float2 row_coeffs(const ScanArgs& a, uint32_t r, const RowRaw x) {
    const uint32_t slot = a.first_slot + r;
    const bool dead = a.deleted != nullptr && ((x.bits >> (slot & 31u)) & 1u);
    float2 ab;
    if (a.metric == L2SQ) ab = make_float2(-2.f, x.xn);
    else if (a.metric == IP) ab = make_float2(-1.f, 0.f);
    else ab = make_float2(x.xn > 0.f ? -rsqrtf(x.xn) : 0.f, 0.f);
    if (r >= a.n || dead) ab = make_float2(0.f, INFINITY);
    return ab;
}

// R: sampled keys a thread keeps (sampling pass only): the insertion network costs 2 (R - 1) instructions per accepted key, and a
// key accepted by ONE of a warp's 32 queries is paid for by all of them
// CL: launched as clusters of two CTAs along x (two query tiles, the same corpus slice). Each CTA fetches HALF of every corpus tile and
// the TMA multicasts it into both CTAs' rings, so a tile crosses L2 -> SM once per pair (a third less L2 traffic at d > 128, half at
// d <= 128 where the queries are resident). A stage is free again when BOTH CTAs' MMAs have read it (the commit arrives on both CTAs'
// `empty` barriers). mapX then has a box of BN / 2 rows. Measured on B200: 3.51 -> 3.42 ms at 1M x 128 x 10k, no change at d = 768 / 1536
// (tensor pipe 88-90 % active either way: those scans were not waiting for the L2).
template <bool EMIT, int R, bool CL = false>
__global__ void __launch_bounds__(threads_of(EMIT), 1)
k_tc_scan(const __grid_constant__ CUtensorMap mapQ, const __grid_constant__ CUtensorMap mapX, const ScanArgs a) {
    extern __shared__ unsigned char smem_raw[];
    // round the base up to 1024 B (128 B swizzle atoms) by pointer arithmetic, so the compiler keeps the shared address space
    unsigned char* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);
    float2* colv = reinterpret_cast<float2*>(smem + (size_t)STAGES * STAGE_BYTES);
    Ctrl* ctrl = reinterpret_cast<Ctrl*>(smem + (size_t)STAGES * STAGE_BYTES + COLV_BYTES);
    const uint32_t smem_s = s32(smem);
    // resident query tile: K blocks 0..kblocks-1 at the front, the ring (corpus tiles only) behind them; otherwise a stage = (A, B)
    const bool res_a = a.kblocks <= (uint32_t)RES_KB && !(a.flags & 1u);
    const uint32_t ring_s = smem_s + (res_a ? (uint32_t)(RES_KB * A_BYTES) : 0u), stage_bytes = res_a ? (uint32_t)B_BYTES : (uint32_t)STAGE_BYTES,
                   b_off = res_a ? 0u : (uint32_t)A_BYTES;
    constexpr int EW = epi_warps(EMIT), W_TMA = EW, W_MMA = EW + 1, W_COEF = EW + 2;
    constexpr int EPI_COLS = BN / (EW / 4);   // columns one epilogue warp scans per tile
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t q0 = blockIdx.x * BM, slice = blockIdx.y;
    // second scan: launched for the largest query count it may see; CTAs past the count the first round left on the device leave at once
    uint32_t nq_eff = a.nq;
    if (a.n_active) { nq_eff = min(*a.n_active, a.active_cap); if ((CL ? (blockIdx.x & ~1u) * BM : q0) >= nq_eff) return; }   // (a pair leaves together)
    const uint32_t crank = CL ? cluster_ctarank() : 0u;
    const uint32_t t_begin = slice * a.tiles_per_slice, t_end = min(a.n_tiles, t_begin + a.tiles_per_slice);
    const uint32_t t_step = EMIT ? 1u : a.sample_stride;
    const uint32_t t_first = EMIT ? t_begin : ((t_begin + a.sample_stride - 1) / a.sample_stride) * a.sample_stride;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(s32(&ctrl->full[i]), 1); mbar_init(s32(&ctrl->empty[i]), CL ? 2 : 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(s32(&ctrl->tfull[i]), 1); mbar_init(s32(&ctrl->tempty[i]), EW); }
        for (int i = 0; i < CB; ++i) { mbar_init(s32(&ctrl->cfull[i]), 1); mbar_init(s32(&ctrl->cempty[i]), EW); }
        mbar_init(s32(&ctrl->afull), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == W_MMA) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&ctrl->tmem_base)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL) cluster_sync_all();      // the peer's barriers exist before anything of ours can arrive on them
    tc_fence_after();
    const uint32_t tmem_base = ctrl->tmem_base;

    if (warp == W_TMA) {
        // ===== TMA producer =====
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            bool t_end_skip = false;
            if (a.flags & 16u) {       // measurement: the MMAs run on whatever the shared memory holds
                if (res_a && t_first < t_end) mbar_arrive(s32(&ctrl->afull));
                for (uint32_t t = t_first; t < t_end; t += t_step)
                    for (uint32_t kb = 0; kb < a.kblocks; ++kb) {
                        mbar_wait(s32(&ctrl->empty[stage]), phase ^ 1u);
                        mbar_arrive(s32(&ctrl->full[stage]));
                        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                    }
                t_end_skip = true;
            }
            if (res_a && t_first < t_end && !t_end_skip) {
                mbar_expect_tx(s32(&ctrl->afull), a.kblocks * A_BYTES);
                for (uint32_t kb = 0; kb < a.kblocks; ++kb) tma_load_2d(smem_s + kb * A_BYTES, &mapQ, s32(&ctrl->afull), (int)(kb * BK), (int)q0);
            }
            for (uint32_t t = t_first; t < t_end && !t_end_skip; t += t_step) {
                for (uint32_t kb = 0; kb < a.kblocks; ++kb) {
                    mbar_wait(s32(&ctrl->empty[stage]), phase ^ 1u);
                    const uint32_t full = s32(&ctrl->full[stage]);
                    mbar_expect_tx(full, res_a ? B_BYTES : STAGE_BYTES);
                    const uint32_t dst = ring_s + stage * stage_bytes;
                    if (!res_a) tma_load_2d(dst, &mapQ, full, (int)(kb * BK), (int)q0);
                    if (CL) tma_load_2d_mc2(dst + b_off + crank * (uint32_t)(B_BYTES / 2), &mapX, full, (int)(kb * BK), (int)(t * BN + crank * (BN / 2)));
                    else tma_load_2d(dst + b_off, &mapX, full, (int)(kb * BK), (int)(t * BN));
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == W_MMA) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            uint32_t stage = 0, phase = 0, it = 0;
            if (res_a && t_first < t_end) mbar_wait(s32(&ctrl->afull), 0u);
            for (uint32_t t = t_first; t < t_end; t += t_step, ++it) {
                const uint32_t buf = it & 1u;
                mbar_wait(s32(&ctrl->tempty[buf]), ((it >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * BN;
                for (uint32_t kb = 0; kb < a.kblocks; ++kb) {
                    mbar_wait(s32(&ctrl->full[stage]), phase);
                    tc_fence_after();
                    const uint64_t da = smem_desc_sw128(res_a ? smem_s + kb * A_BYTES : ring_s + stage * stage_bytes);
                    const uint64_t db = smem_desc_sw128(ring_s + stage * stage_bytes + b_off);
#ifndef TC_DEBUG_SKIP_MMA
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
#else
                    if (kb == 0xffffffffu) for (int k = 0; k < 1; ++k)
#endif
                        umma_tf32(d_tmem, da + (uint64_t)(k * UMMA_K * 4 >> 4), db + (uint64_t)(k * UMMA_K * 4 >> 4), IDESC_TF32, (kb | (uint32_t)k) != 0u);
                    if (CL) umma_commit_mc2(s32(&ctrl->empty[stage])); else umma_commit(s32(&ctrl->empty[stage]));      // frees the smem stage once these MMAs are done
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit(s32(&ctrl->tfull[buf]));              // accumulator ready for the epilogue
            }
        }
    } else if (warp == W_COEF) {
        // ===== row coefficients (a_j, b_j) of the tiles, through a ring of CB buffers: the global loads of CG tiles are in flight
        // together, and the warp runs ahead of the epilogue by up to CB tiles =====
        uint32_t it = 0;
        const int cg = (a.flags & 2u) ? 1 : CG;
        for (uint32_t t = t_first; t < t_end; t += cg * t_step) {
            RowRaw raw[CG][BN / 32];
#pragma unroll
            for (int gi = 0; gi < CG; ++gi) {
                const uint32_t tg = min(t + (gi < cg ? gi : 0) * t_step, t_end - 1u);      // (rounds past the end re-read a valid tile and drop it)
#pragma unroll
                for (int j = 0; j < BN / 32; ++j) raw[gi][j] = row_load(a, tg * BN + j * 32 + lane);
            }
#pragma unroll
            for (int gi = 0; gi < CG; ++gi) {
                if (t + gi * t_step < t_end && gi < cg) {
                    const uint32_t cb = it & (CB - 1);
                    mbar_wait(s32(&ctrl->cempty[cb]), ((it / CB) & 1u) ^ 1u);       // the epilogue is done with colv[cb]
#pragma unroll
                    for (int j = 0; j < BN / 32; ++j) colv[cb * BN + j * 32 + lane] = row_coeffs(a, (t + gi * t_step) * BN + j * 32 + lane, raw[gi][j]);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(s32(&ctrl->cfull[cb]));
                    ++it;
                }
            }
        }
    } else {
        // ===== epilogue: thread = one query row of the accumulator =====
        // warp w owns TMEM lanes 32 (w % 4) .. +31 (hardware rule) and the column half w / 4
        const uint32_t row = (uint32_t)(warp & 3) * 32 + lane, half = (uint32_t)warp >> 2, q = q0 + row;
        const bool qok = q < nq_eff;
        float tau = -INFINITY;
        if (EMIT && qok) tau = a.tau[q];
        float top[R];
#pragma unroll
        for (int i = 0; i < R; ++i) top[i] = INFINITY;
        uint32_t pend_hits = 0, pend_pos = 0, pend_base = 0;
        auto flush_pending = [&]() {
            while (pend_hits) {
                const uint32_t i = (uint32_t)__ffs(pend_hits) - 1u; pend_hits &= pend_hits - 1u;
                if (pend_pos < a.cap) a.cand[(size_t)q * a.cap + pend_pos] = pend_base + i;
                ++pend_pos;
            }
        };
        uint32_t it = 0;
        for (uint32_t t = t_first; t < t_end; t += t_step, ++it) {
            const uint32_t buf = it & 1u, cb = it & (CB - 1);
            mbar_wait(s32(&ctrl->cfull[cb]), (it / CB) & 1u);
            mbar_wait(s32(&ctrl->tfull[buf]), (it >> 1) & 1u);
            tc_fence_after();
            const float2* cv = colv + cb * BN + half * EPI_COLS;
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + buf * BN + half * EPI_COLS;
            // Hot loop: key = a * dot + b and running minima over groups of 8 columns, branch-free; the (rare) groups whose minimum
            // passes the bound are re-scanned element by element. The next 32 columns are in flight while these are processed.
            const float4* cv4 = reinterpret_cast<const float4*>(cv);
            auto process = [&](const float (&v)[32], int c) {
                float m[4];
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    float mg = INFINITY;
#pragma unroll
                    for (int i = g * 8; i < g * 8 + 8; i += 2) {
                        const float4 ab = cv4[(c * 32 + i) >> 1];
                        mg = fminf(mg, fminf(fmaf(ab.x, v[i], ab.y), fmaf(ab.z, v[i + 1], ab.w)));
                    }
                    m[g] = mg;
                }
                const float mn = fminf(fminf(m[0], m[1]), fminf(m[2], m[3]));
                if (EMIT) {
                    if (mn <= tau) {
                        uint32_t hits = 0;
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            if (m[g] <= tau) {
#pragma unroll
                                for (int i = g * 8; i < g * 8 + 8; ++i) {   // unrolled: v[] stays in registers
                                    const float2 ab = cv[c * 32 + i];
                                    if (fmaf(ab.x, v[i], ab.y) <= tau) hits |= 1u << i;
                                }
                            }
                        }
                        // one atomic per chunk with candidates, and nobody waits for it: the ids are written when the next such
                        // chunk comes by (or at the end), long after the position has come back
                        flush_pending();
                        pend_pos = atomicAdd(&a.cnt[q], (uint32_t)__popc(hits));
                        pend_hits = hits; pend_base = a.first_slot + t * BN + half * EPI_COLS + c * 32;
                        if (a.flags & 4u) flush_pending();
                    }
                } else if (mn < top[R - 1]) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        if (m[g] < top[R - 1]) {
#pragma unroll
                            for (int i = g * 8; i < g * 8 + 8; ++i) {
                                const float2 ab = cv[c * 32 + i];
                                const float key = fmaf(ab.x, v[i], ab.y);
                                if (key < top[R - 1]) {
                                    top[R - 1] = key;
#pragma unroll
                                    for (int j = R - 1; j > 0; --j) {
                                        const float lo = fminf(top[j - 1], top[j]), hi = fmaxf(top[j - 1], top[j]);
                                        top[j - 1] = lo; top[j] = hi;
                                    }
                                }
                            }
                        }
                    }
                }
            };
#ifndef TC_DEBUG_SKIP_EPI
            float va[32], vb[32];
            if (!(a.flags & 8u)) tmem_ld32_issue(taddr, va);
#pragma unroll 1
            for (int c = 0; c < EPI_COLS / 32 && !(a.flags & 8u); c += 2) {
                tmem_wait();
                tmem_ld32_issue(taddr + (c + 1) * 32, vb);
                process(va, c);
                tmem_wait();
                if (c + 2 < EPI_COLS / 32) tmem_ld32_issue(taddr + (c + 2) * 32, va);
                process(vb, c + 1);
            }
#endif
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { mbar_arrive(s32(&ctrl->tempty[buf])); mbar_arrive(s32(&ctrl->cempty[cb])); }
        }
        if (EMIT) flush_pending();
        if (!EMIT && qok) {
            float* o = a.samp + ((size_t)(slice * (EPI_WARPS / 4) + half) * a.nq_pad + q) * R_MAX;
#pragma unroll
            for (int i = 0; i < R_MAX; ++i) o[i] = i < R ? top[i] : INFINITY;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CL) cluster_sync_all();      // the peer may still be arriving on this CTA's barriers
    if (warp == W_MMA) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}


// Worst-case |approximate distance - exact f32 distance| for one query (see the file header).
__device__ __forceinline__ float tc_eps(int metric, float qn, float max_xn, uint32_t dpad) {
    const float e_dot = 0.001953125f /* 2^-9: both operands truncated to TF32 */ + (float)dpad * 4.76837158e-7f /* d 2^-21: accumulation */;
    const float e_f32 = (float)(dpad + 4) * 1.1920929e-7f;   // (d + 4) 2^-23: the two f32 chains
    const float qx = sqrtf(qn) * sqrtf(max_xn);
    float e;
    if (metric == L2SQ) e = 2.f * qx * e_dot + (qn + max_xn) * e_f32 * 2.f;
    else if (metric == IP) e = qx * (e_dot + e_f32);
    else e = e_dot + e_f32 + 1e-6f;
    return e * 1.25f;   // head-room for the rounding of this formula itself
}

// tau[q] = r-th smallest of the keys sampled for q over all slices. One block per query: the samples are staged in
// shared memory once, then r rounds of block-wide "smallest key after the previous pick" ((value, index) order, so
// equal keys count separately). With a slice per CTA and several waves of CTAs there are thousands of samples per
// query; the former one-warp-per-query form re-read them from global memory r times.
constexpr int THR_THREADS = 128;
__global__ void __launch_bounds__(THR_THREADS) k_tc_threshold(const float* __restrict__ samp, uint32_t slices, uint32_t nq, uint32_t nq_pad, uint32_t r,
                                                             uint32_t r0, float* __restrict__ tau, const float* __restrict__ qn,
                                                             const uint32_t* __restrict__ max_xn_bits, int metric, uint32_t dpad, uint32_t stride, uint32_t cap, int widen) {
    extern __shared__ __align__(16) unsigned char thr_sm[];
    float* v = reinterpret_cast<float*>(thr_sm);
    __shared__ float wv[THR_THREADS / 32]; __shared__ uint32_t wi[THR_THREADS / 32];
    const uint32_t q = blockIdx.x, t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    const uint32_t total = slices * R_MAX;
    for (uint32_t idx = t; idx < total; idx += THR_THREADS) v[idx] = samp[((size_t)(idx / R_MAX) * nq_pad + q) * R_MAX + (idx % R_MAX)];
    __syncthreads();
    float pv = -INFINITY, p0 = -INFINITY; uint32_t pi = 0; bool have = false;
    for (uint32_t round = 0; round < r; ++round) {
        float bv = INFINITY; uint32_t bi = NONE;
        for (uint32_t idx = t; idx < total; idx += THR_THREADS) {
            const float x = v[idx];
            const bool after = !have || x > pv || (x == pv && idx > pi);
            if (after && (x < bv || (x == bv && idx < bi))) { bv = x; bi = idx; }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (oi != NONE && (bi == NONE || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
        }
        if (lane == 0) { wv[warp] = bv; wi[warp] = bi; }
        __syncthreads();
        bv = wv[0]; bi = wi[0];
#pragma unroll
        for (int w = 1; w < THR_THREADS / 32; ++w) {
            const float ov = wv[w]; const uint32_t oi = wi[w];
            if (oi != NONE && (bi == NONE || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
        }
        __syncthreads();
        if (bi == NONE) { pv = INFINITY; break; }   // fewer than r samples: no usable threshold
        pv = bv; pi = bi; have = true;
        if (round + 1 == r0) p0 = pv;
    }
    // Widening. The proof in k_tc_select asks for d_k < dist(tau) - eps. p0, the sampled key of rank r0 ~ k / stride, estimates
    // the k-th neighbour's key from above; when tau - p0 < 2 eps (dense neighbourhoods: 5M x 1536 clustered, k = 100 has
    // hundreds of rows within eps of the k-th, two thirds of the queries failed the proof and paid a second full scan, 9.9 ms
    // instead of 5.5 ms at B = 64) the threshold is raised by 2.5 eps: whenever k live rows have an approximate key <= the
    // sampled tau, their exact distances are <= dist(tau) + eps < dist(tau + 2.5 eps) - eps and the proof holds. Not when the
    // sample says the wider threshold would overflow the candidate buffer (12.5M x 96 L2: the bound is loose there, the
    // unwidened proof succeeds for 99.9 % of the queries and widening made 5 % overflow).
    float e = tc_eps(metric, qn[q], __uint_as_float(*max_xn_bits), dpad);
    if (metric == COS) e *= sqrtf(qn[q]);          // dist = 1 + key / |q|
    const float tw = pv + 2.5f * e;
    uint32_t below = 0;
    for (uint32_t idx = t; idx < total; idx += THR_THREADS) below += v[idx] <= tw ? 1u : 0u;
    below = __reduce_add_sync(0xffffffffu, below);
    if (lane == 0) wi[warp] = below;
    __syncthreads();
    if (t == 0) {
        below = 0;
        for (int w = 0; w < THR_THREADS / 32; ++w) below += wi[w];
        const bool dense = (pv - p0) < 2.f * e, fits = (uint64_t)below * stride * 2u <= (uint64_t)cap;
        tau[q] = (widen && dense && fits && isfinite(pv) && isfinite(tw)) ? tw : pv;
    }
}

__global__ void k_tc_max_norm(const float* __restrict__ sqn, uint32_t n, uint32_t* out_bits) {
    float m = 0.f;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = fmaxf(m, sqn[i]);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));   // non-negative floats order like their bits
}

struct RerankArgs {
    const float* vecs; const float* sqn; const uint64_t* slot_key; const uint32_t* deleted;
    uint32_t first_slot, n_slots, dpad;
    const float* q; size_t q_stride; const float* qn; uint32_t nq;     // the queries of this round (local index)
    const uint32_t* qmap;                                              // local -> caller's query index (null = identity)
    int metric; uint32_t k, kp, cap;
    uint32_t* cand; float* cand_d; const uint32_t* cnt; const float* tau; const uint32_t* max_xn_bits;
    uint64_t* out_keys; uint32_t* out_slots; float* out_dist; uint32_t out_ld;   // out_keys and/or out_slots, rows of out_ld entries
    uint32_t self_first, self_step;                                    // query g is slot self_first + g * self_step (excluded); 0 = none
    // round 1 only: queries whose proof failed but whose k-th distance is known go to a second scan with
    // tau2 = key(dk + 2 eps) (a superset of the true top-k by construction)
    uint32_t* n_r2; uint32_t* r2_ids; float* r2_tau; float* r2_qn;
    uint32_t* n_fb; uint32_t* fb_ids;                                  // queries only the FP32 fallback can answer
    const uint32_t* n_active;                                          // second round: number of queries, on the device (null = nq)
    uint32_t r2_cap;                                                   // second round: at most this many queries (the rest goes to the fallback)
};
__device__ __forceinline__ uint32_t rerank_nq(const RerankArgs& a) { return a.n_active ? min(*a.n_active, a.nq) : a.nq; }

// Exact distances of the emitted candidates: grid (cap / 128, nq), one warp per 32 candidates. Each lane owns one
// candidate and runs the oracle's sequential fmaf chain over the dimensions; rows are staged through a
// transposed shared tile so that the global reads are 128 B coalesced.
__global__ void __launch_bounds__(128) k_tc_exact_dist(RerankArgs a) {
    extern __shared__ __align__(16) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t q = blockIdx.x, cblk = blockIdx.y;   // queries on x: batches above 65535 queries
    if (q >= rerank_nq(a)) return;
    const uint32_t cnt = min(a.cnt[q], a.cap);
    if (cblk * 128u >= cnt) return;                          // whole block past the end
    float* qs = reinterpret_cast<float*>(sm);
    float* tile = qs + a.dpad + (size_t)warp * 32 * 33;
    for (uint32_t c = threadIdx.x; c < a.dpad; c += blockDim.x) qs[c] = a.q[(size_t)q * a.q_stride + c];
    __syncthreads();
    const uint32_t b0 = (cblk * 4 + warp) * 32;
    if (b0 >= cnt) return;
    uint32_t slot = b0 + lane < cnt ? a.cand[(size_t)q * a.cap + b0 + lane] : NONE;
    if (slot != NONE && a.deleted && ((a.deleted[slot >> 5] >> (slot & 31u)) & 1u)) slot = NONE;
    if (a.self_step && slot == a.self_first + (a.qmap ? a.qmap[q] : q) * a.self_step) slot = NONE;
    float acc = 0.f;
    for (uint32_t kc = 0; kc < a.dpad; kc += 32) {
        // 32 rows x 32 floats = 256 float4: lane l fetches float4 (it * 32 + l): row = idx / 8, quad = idx % 8
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * 32 + lane, r = idx >> 3, c4 = (idx & 7) * 4;
            const uint32_t sj = __shfl_sync(0xffffffffu, slot, r);
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sj != NONE && kc + c4 < a.dpad) x = __ldg(reinterpret_cast<const float4*>(a.vecs + (size_t)(sj - a.first_slot) * a.dpad + kc + c4));
            float* t = tile + r * 33 + c4;
            t[0] = x.x; t[1] = x.y; t[2] = x.z; t[3] = x.w;
        }
        __syncwarp();
        const uint32_t kend = min(32u, a.dpad - kc);
        if (a.metric == L2SQ) { for (uint32_t i = 0; i < kend; ++i) { const float d = qs[kc + i] - tile[lane * 33 + i]; acc = fmaf(d, d, acc); } }
        else { for (uint32_t i = 0; i < kend; ++i) acc = fmaf(qs[kc + i], tile[lane * 33 + i], acc); }
        __syncwarp();
    }
    float d = INFINITY;
    if (slot != NONE) d = finish_metric(a.metric, acc, a.qn[q], a.metric == COS ? a.sqn[slot - a.first_slot] : 0.f);
    if (b0 + lane < cnt) { a.cand_d[(size_t)q * a.cap + b0 + lane] = d; a.cand[(size_t)q * a.cap + b0 + lane] = slot; }
}

// One warp per query: top-k of the candidates by (distance, key), then the proof that no row outside the
// candidate set can belong to the top-k (rows that were not emitted have key > tau, i.e. exact distance >
// dist(tau) - eps).
template <class Less>
__global__ void __launch_bounds__(128) k_tc_select(RerankArgs a, Less less) {
    extern __shared__ __align__(16) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    const uint32_t q = blockIdx.x * wpb + warp;
    if (q >= rerank_nq(a)) return;
    unsigned char* base = sm + (size_t)warp * ((size_t)a.kp * 16 + 256);
    SList L;
    L.ld = reinterpret_cast<float*>(base); L.ld2 = L.ld + a.kp; L.li = reinterpret_cast<uint32_t*>(L.ld2 + a.kp); L.li2 = L.li + a.kp;
    uint32_t* scr_id = L.li2 + a.kp; float* scr_d = reinterpret_cast<float*>(scr_id + 32);
    L.cap = (int)a.k; L.capp = (int)a.kp;
    slist_init(L);
    const uint32_t total = a.cnt[q], cnt = min(total, a.cap);
    for (uint32_t b0 = 0; b0 < cnt; b0 += 32) {
        const uint32_t slot = b0 + lane < cnt ? a.cand[(size_t)q * a.cap + b0 + lane] : NONE;
        const float d = slot != NONE ? a.cand_d[(size_t)q * a.cap + b0 + lane] : INFINITY;
        slist_merge(L, d, slot, scr_id, scr_d, less);
    }
    const float qn = a.qn[q], tau = a.tau[q], max_xn = __uint_as_float(*a.max_xn_bits);
    float tau_dist;
    if (a.metric == L2SQ) tau_dist = qn + tau;
    else if (a.metric == IP) tau_dist = 1.f + tau;
    else tau_dist = qn > 0.f ? 1.f + tau * rsqrtf(qn) : -INFINITY;
    const float eps = tc_eps(a.metric, qn, max_xn, a.dpad);
    const float dk = L.ld[a.k - 1];
    const bool full = L.li[a.k - 1] != NONE, fits = total <= a.cap;
    const uint32_t gq = a.qmap ? a.qmap[q] : q;
    if (fits && full && dk < tau_dist - eps) {
        for (uint32_t i = lane; i < a.out_ld; i += 32) {
            const uint32_t id = i < a.k ? L.li[i] : NONE;
            if (a.out_slots) a.out_slots[(size_t)gq * a.out_ld + i] = id;
            if (a.out_keys) a.out_keys[(size_t)gq * a.out_ld + i] = id == NONE ? 0xFFFFFFFFFFFFFFFFull : a.slot_key[id];
            if (a.out_dist) a.out_dist[(size_t)gq * a.out_ld + i] = i < a.k ? L.ld[i] : INFINITY;
        }
    } else if (lane == 0) {
        const float t2d = dk + 2.f * eps;
        float tau2;
        if (a.metric == L2SQ) tau2 = t2d - qn;
        else if (a.metric == IP) tau2 = t2d - 1.f;
        else tau2 = (t2d - 1.f) * sqrtf(qn);
        // (a truncated candidate set still bounds the k-th distance from above, so overflow is no obstacle here)
        uint32_t i = 0xFFFFFFFFu;
        if (a.n_r2 && full && (a.metric != COS || qn > 0.f) && isfinite(tau2)) i = atomicAdd(a.n_r2, 1u);
        if (i < a.r2_cap) { a.r2_ids[i] = gq; a.r2_tau[i] = tau2; a.r2_qn[i] = qn; }
        else a.fb_ids[atomicAdd(a.n_fb, 1u)] = gq;
    }
}

// gather / scatter for the queries that go to the FP32 kernel
__global__ void k_tc_gather_queries(const float* __restrict__ q, size_t q_stride, uint32_t dpad, const uint32_t* __restrict__ flagged, const uint32_t* __restrict__ n_dev, uint32_t cap,
                                    float* __restrict__ out) {
    const size_t total = (size_t)min(*n_dev, cap) * dpad;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
        out[i] = q[(size_t)flagged[i / dpad] * q_stride + (i % dpad)];
}
// The queries neither round could prove (fewer than k live candidates, or a candidate list that overflowed twice: thousands of
// exact ties): a plain FP32 scan with the oracle's sequential fmaf chain, one BLOCK per flagged query, driven by the count the
// select kernels left on the device - launched unconditionally, it leaves at once when that count is zero, so the host never has
// to read a counter back (no stream synchronisation anywhere in the exact path). Tiles of 256 rows go through a transposed shared
// tile (coalesced loads, thread-per-row chains); rows that can still enter the top-k are merged into the block's sorted list.
template <class Less>
__global__ void __launch_bounds__(256) k_tc_fallback(RerankArgs a, uint32_t n_rows, Less less) {
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ uint32_t nbuf; __shared__ float worst;
    const uint32_t nf = *a.n_fb;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float* qs = reinterpret_cast<float*>(sm);
    float* tile = qs + a.dpad;
    float* bd = tile + 256 * 33;
    uint32_t* bi = reinterpret_cast<uint32_t*>(bd + 256);
    SList L;
    L.ld = reinterpret_cast<float*>(bi + 256); L.ld2 = L.ld + a.kp; L.li = reinterpret_cast<uint32_t*>(L.ld2 + a.kp); L.li2 = L.li + a.kp;
    uint32_t* scr_id = L.li2 + a.kp; float* scr_d = reinterpret_cast<float*>(scr_id + 32);
    L.cap = (int)a.k; L.capp = (int)a.kp;
    for (uint32_t f = blockIdx.x; f < nf; f += gridDim.x) {
        const uint32_t gq = a.fb_ids[f];
        __syncthreads();
        for (uint32_t c = tid; c < a.dpad; c += blockDim.x) qs[c] = a.q[(size_t)gq * a.q_stride + c];
        if (warp == 0) slist_init(L);
        if (tid == 0) { nbuf = 0; worst = INFINITY; }
        __syncthreads();
        const float qn = a.qn[gq];
        const uint32_t self = a.self_step ? a.self_first + gq * a.self_step : NONE;
        for (uint32_t r0 = 0; r0 < n_rows; r0 += 256) {
            const uint32_t row = r0 + tid, slot = a.first_slot + row;
            bool valid = row < n_rows && slot != self;
            if (valid && a.deleted && ((a.deleted[slot >> 5] >> (slot & 31u)) & 1u)) valid = false;
            float acc = 0.f;
            for (uint32_t kc = 0; kc < a.dpad; kc += 32) {
                __syncthreads();
#pragma unroll 4
                for (int it = 0; it < 32; ++it) {
                    const int idx = it * 256 + tid, rr = idx >> 5, cc = idx & 31;
                    const uint32_t r = r0 + rr;
                    tile[rr * 33 + cc] = (r < n_rows && kc + cc < a.dpad) ? __ldg(a.vecs + (size_t)r * a.dpad + kc + cc) : 0.f;
                }
                __syncthreads();
                const uint32_t kend = min(32u, a.dpad - kc);
                if (a.metric == L2SQ) { for (uint32_t i = 0; i < kend; ++i) { const float d = qs[kc + i] - tile[tid * 33 + i]; acc = fmaf(d, d, acc); } }
                else { for (uint32_t i = 0; i < kend; ++i) acc = fmaf(qs[kc + i], tile[tid * 33 + i], acc); }
            }
            const float d = valid ? finish_metric(a.metric, acc, qn, a.metric == COS ? a.sqn[row] : 0.f) : INFINITY;
            const bool ok = valid && d <= worst;                       // ties are admitted: the key order decides in the merge
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            uint32_t base = 0;
            if (lane == 0 && m) base = atomicAdd(&nbuf, (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (ok) { const uint32_t pos = base + __popc(m & ((1u << lane) - 1u)); bd[pos] = d; bi[pos] = slot; }
            __syncthreads();
            const uint32_t nb = nbuf;
            if (nb && warp == 0) {
                for (uint32_t b0 = 0; b0 < nb; b0 += 32) {
                    const bool on = b0 + lane < nb;
                    slist_merge(L, on ? bd[b0 + lane] : INFINITY, on ? bi[b0 + lane] : NONE, scr_id, scr_d, less);
                }
                if (lane == 0) { worst = L.ld[a.k - 1]; nbuf = 0; }
            }
            __syncthreads();
        }
        if (warp == 0) {
            for (uint32_t i = lane; i < a.out_ld; i += 32) {
                const uint32_t id = i < a.k ? L.li[i] : NONE;
                if (a.out_slots) a.out_slots[(size_t)gq * a.out_ld + i] = id;
                if (a.out_keys) a.out_keys[(size_t)gq * a.out_ld + i] = id == NONE ? 0xFFFFFFFFFFFFFFFFull : a.slot_key[id];
                if (a.out_dist) a.out_dist[(size_t)gq * a.out_ld + i] = (i < a.k && id != NONE) ? L.ld[i] : INFINITY;
            }
        }
    }
}

// second_scan += number of queries that needed the second round, fallback += number the FP32 fallback answered (vs_get_stats)
__global__ void k_tc_stats(const uint32_t* __restrict__ ctr, unsigned long long* __restrict__ dst) {
    if (threadIdx.x == 0 && blockIdx.x == 0) { dst[0] += ctr[0]; dst[1] += ctr[2]; }
}

// ---- host side ---------------------------------------------------------------------------------
// rows x dpad f32 matrix, row-major; box = box_rows x 32 floats, 128 B swizzle, zero fill outside
static bool make_map(CUtensorMap* m, const float* base, uint64_t rows, uint32_t dpad, size_t row_stride_floats, uint32_t box_rows) {
    return make_tmap_f32(m, base, rows, dpad, row_stride_floats, BK, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

struct Plan { uint32_t nqt, nqt_launch, nq_pad, n_tiles, slices, tiles_per_slice, s_slices, s_tiles_per_slice, kblocks, stride, r, cap, kp; bool ok, cluster; };

// r-th smallest of a 1/stride sample as threshold: expected stride * r rows pass, P(fewer than k pass) < 1e-3
static uint32_t sample_rank(uint32_t k, uint32_t stride) {
    const float lam = (float)k / stride;
    return (uint32_t)std::ceil(lam + 4.f * std::sqrt(lam) + 3.5f);
}

// Slices per query tile in [lo, hi] whose grid fills its last wave best (one CTA per SM, CTAs of equal length): 79 query tiles x 15
// slices = 148 x 8 + 1 CTAs ran a ninth wave for a single CTA (11 % of the scan). Ties go to the finer slicing.
static uint32_t pick_slices(uint32_t nqt, uint32_t lo, uint32_t hi, int sm_count) {
    uint32_t best = hi; double best_eff = -1.0;
    for (uint32_t s = hi; s >= lo && s >= 1; --s) {
        const double waves = (double)nqt * s / sm_count, eff = waves / std::ceil(waves);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
        if (s == 1) break;
    }
    return best;
}

static Plan make_plan(uint32_t nq, uint32_t n, uint32_t k, uint32_t dpad, int sm_count) {
    Plan p;
    p.nqt = (nq + BM - 1) / BM; p.nq_pad = p.nqt * BM;
    // full scan as 2-CTA clusters (the corpus tile multicast to both): from two query tiles on; an odd count gets one idle partner
    static const int cl_env = [] { const char* e = getenv("VSANN_TC_CLUSTER"); return e ? atoi(e) : 1; }();
    p.cluster = cl_env != 0 && p.nqt >= 2 && (p.nqt % 2 == 0 || p.nqt >= 32);   // (an idle partner is 1 / nqt of wasted MMAs)
    p.nqt_launch = p.cluster ? (p.nqt + 1u) & ~1u : p.nqt;
    p.n_tiles = (n + BN - 1) / BN;
    p.kblocks = (dpad + BK - 1) / BK;
    uint32_t want = (8u * (uint32_t)sm_count + p.nqt_launch - 1) / p.nqt_launch;          // ~8 waves of CTAs
    want = std::min<uint32_t>(want, std::max<uint32_t>(1u, p.n_tiles / 8));   // >= 8 tiles per CTA
    // (a cap of 64 slices left 64 of 148 SMs busy for a single query tile: 2.3 TB/s at 5M x 1536, B <= 128)
    want = std::max<uint32_t>(1u, std::min<uint32_t>(want, 1024u));
    want = pick_slices(p.nqt_launch, std::max<uint32_t>(1u, (want + 1) / 2), want, sm_count);
    p.tiles_per_slice = std::max<uint32_t>(1u, (p.n_tiles + want - 1) / want);
    p.slices = (p.n_tiles + p.tiles_per_slice - 1) / p.tiles_per_slice;
    // sampling stride: 32 on large corpora (the sample pass is the slow one per tile), denser on small ones so
    // that fewer rows pass the threshold; never so dense that the rank exceeds the R_MAX keys a thread keeps
    uint32_t smin = 1; while (smin < SAMPLE_STRIDE && sample_rank(k, smin) > R_MAX) ++smin;
    p.stride = p.n_tiles >= 512 ? SAMPLE_STRIDE : std::min<uint32_t>(SAMPLE_STRIDE, std::max<uint32_t>(smin, p.n_tiles / 16));
    p.ok = sample_rank(k, p.stride) <= R_MAX && p.n_tiles >= 2 * p.stride;
    p.r = std::min<uint32_t>(R_MAX, sample_rank(k, p.stride));
    // The sampling pass has its own, coarser slicing (~2 waves of CTAs, at least one sampled tile each): a thread that sees more
    // keys accepts a smaller share of them (R ln(m / R) of m), and with 8 tiles per CTA 85 % of the sampled keys went through
    // the insertion network of all 32 lanes (0.67 ms for 1/32 of the work of the 3.8 ms full scan)
    uint32_t want_s = (2u * (uint32_t)sm_count + p.nqt - 1) / p.nqt;
    const uint32_t cap_s = std::max<uint32_t>(1u, std::min<uint32_t>(p.slices, std::max<uint32_t>(1u, p.n_tiles / p.stride)));
    want_s = pick_slices(p.nqt, std::max<uint32_t>(1u, std::min<uint32_t>(cap_s, (want_s + 1) / 2)), std::max<uint32_t>(1u, std::min<uint32_t>(cap_s, 2 * want_s)), sm_count);
    p.s_tiles_per_slice = std::max<uint32_t>(1u, (p.n_tiles + want_s - 1) / want_s);
    p.s_slices = (p.n_tiles + p.s_tiles_per_slice - 1) / p.s_tiles_per_slice;
    uint32_t cap = 1; while (cap < (k >= 32 ? 8u : 4u) * p.r * p.stride) cap <<= 1;   // k >= 32: room for a widened threshold in dense data
    p.cap = cap;
    p.kp = (k + 31u) & ~31u;
    return p;
}

static inline size_t al256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace tc

bool exact_tc_supported(const ExactCorpus& c, const ExactQueries& q, uint32_t k) {
    const char* env = getenv("VSANN_EXACT_TC");   // 0 = FP32 kernel only, 1 = tensor path whenever it applies, default: by size
    const int mode = env ? atoi(env) : -1;
    if (mode == 0) return false;
    if (c.row_step != 1 || k == 0 || k > 128 || c.n == 0 || q.nq == 0) return false;
    if (q.stride != q.dpad || c.row_stride != q.dpad) return false;
    if (!tc::make_plan(q.nq, c.n, k + (q.self_step ? 1u : 0u), q.dpad, 148).ok) return false;
    if (mode == 1) return true;
    return (uint64_t)c.n * q.nq >= (1ull << 24);
}

namespace tc {
struct Scratch {
    float *qn, *tau, *samp, *cand_d, *r2_tau, *r2_qn, *gq;
    uint32_t *cnt, *cand, *ctr /* [0] n_r2, [1] max |x|^2 bits, [2] n_fb */, *r2_ids, *fb_ids;
};
static size_t carve(const Plan& p, uint32_t nq, uint32_t dpad, unsigned char* base, Scratch* s) {
    size_t off = 0;
    auto take = [&](size_t bytes) { unsigned char* r = base ? base + off : nullptr; off += al256(bytes); return r; };
    Scratch t{};
    t.qn = (float*)take((size_t)nq * 4); t.tau = (float*)take((size_t)nq * 4); t.cnt = (uint32_t*)take((size_t)nq * 4);
    t.samp = (float*)take((size_t)p.s_slices * (EPI_WARPS / 4) * p.nq_pad * R_MAX * 4);
    t.cand = (uint32_t*)take((size_t)nq * p.cap * 4); t.cand_d = (float*)take((size_t)nq * p.cap * 4);
    t.ctr = (uint32_t*)take(16);
    t.r2_ids = (uint32_t*)take((size_t)nq * 4); t.r2_tau = (float*)take((size_t)nq * 4); t.r2_qn = (float*)take((size_t)nq * 4);
    t.fb_ids = (uint32_t*)take((size_t)nq * 4);
    t.gq = (float*)take((size_t)nq * dpad * 4);
    if (s) *s = t;
    return off + 256;
}
}  // namespace tc

size_t exact_tc_scratch_bytes(uint32_t nq, uint32_t n, uint32_t k, uint32_t dpad, int sm_count) {
    return tc::carve(tc::make_plan(nq, n, k, dpad, sm_count), nq, dpad, nullptr, nullptr);
}

// Exact top-k of every query; results are those of launch_exact (FP32 kernel) bit for bit.
// fb_scratch: scratch of the FP32 kernel, sized for nq queries.
cudaError_t launch_exact_tc(const ExactCorpus& c, const ExactQueries& q, int metric, uint32_t k, uint32_t out_ld,
                            uint32_t* out_slots, uint64_t* out_keys, float* out_dist,
                            const ExactScratch& scratch, const ExactScratch& fb_scratch, const DeviceInfo& dev, cudaStream_t st, TcStats* stats) {
    using namespace tc;
    if (q.nq == 0) return cudaSuccess;
    if (out_ld < k || (out_keys && !c.slot_key)) return cudaErrorInvalidValue;
    const uint32_t kplan = k + (q.self_step ? 1u : 0u);     // the query itself is among the candidates and dropped later
    const Plan p = make_plan(q.nq, c.n, kplan, q.dpad, dev.sm_count);
    if (scratch.bytes < exact_tc_scratch_bytes(q.nq, c.n, kplan, q.dpad, dev.sm_count)) return cudaErrorInvalidValue;
    Scratch S; carve(p, q.nq, q.dpad, (unsigned char*)scratch.ptr, &S);
    CUtensorMap mapQ, mapX;
    if (!make_map(&mapQ, q.q, q.nq, q.dpad, q.stride, BM) || !make_map(&mapX, c.vecs, c.n, q.dpad, c.row_stride, BN)) return cudaErrorNotSupported;
    cudaError_t e;
#define TC_TRY(x) do { if ((e = (x)) != cudaSuccess) return e; } while (0)
    TC_TRY(cudaFuncSetAttribute(k_tc_scan<false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    TC_TRY(cudaFuncSetAttribute(k_tc_scan<false, R_MAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    TC_TRY(cudaFuncSetAttribute(k_tc_scan<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    TC_TRY((cudaFuncSetAttribute(k_tc_scan<true, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES)));
    const size_t dsmem = (size_t)q.dpad * 4 + 4 * 32 * 33 * 4, ssmem = 4 * ((size_t)p.kp * 16 + 256);
    if (dsmem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
    TC_TRY(cudaFuncSetAttribute(k_tc_exact_dist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsmem));
    TC_TRY(cudaMemsetAsync(S.cnt, 0, (size_t)q.nq * 4, st));
    TC_TRY(cudaMemsetAsync(S.ctr, 0, 16, st));
    k_query_norms_tc<<<(q.nq + 127) / 128, 128, 0, st>>>(q.q, q.stride, q.dpad, q.nq, S.qn);
    k_tc_max_norm<<<dev.sm_count, 256, 0, st>>>(c.sqn, c.n, S.ctr + 1);

    ScanArgs a{};
    a.sqn = c.sqn; a.deleted = c.deleted; a.first_slot = c.first_slot; a.n = c.n; a.nq = q.nq; a.nq_pad = p.nq_pad; a.kblocks = p.kblocks;
    a.metric = metric; a.n_tiles = p.n_tiles; a.tiles_per_slice = p.tiles_per_slice; a.sample_stride = p.stride;
    a.samp = S.samp; a.tau = S.tau; a.cand = S.cand; a.cnt = S.cnt; a.cap = p.cap;
    if (const char* fe = getenv("VSANN_TC_FLAGS")) a.flags = (uint32_t)atoi(fe);
    const dim3 grid(p.nqt, p.slices);
    {
        ScanArgs as = a; as.tiles_per_slice = p.s_tiles_per_slice;
        const dim3 sgrid(p.nqt, p.s_slices);
        if (p.r <= 8) k_tc_scan<false, 8><<<sgrid, threads_of(false), SMEM_BYTES, st>>>(mapQ, mapX, as);
        else k_tc_scan<false, R_MAX><<<sgrid, threads_of(false), SMEM_BYTES, st>>>(mapQ, mapX, as);
    }
    const size_t thr_smem = (size_t)p.s_slices * (EPI_WARPS / 4) * R_MAX * 4;
    if (thr_smem > 48 * 1024) TC_TRY(cudaFuncSetAttribute(k_tc_threshold, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)thr_smem));
    // sample rank whose key estimates the k-th neighbour's from above: ceil(k / stride) (rank 1 when the sample is sparser than k)
    const uint32_t r0 = std::max<uint32_t>(1u, std::min<uint32_t>(p.r > 1 ? p.r - 1 : 1u, (kplan + p.stride - 1) / p.stride));
    const char* we = getenv("VSANN_TC_WIDEN");   // 0 = sampled threshold as is (A/B switch)
    k_tc_threshold<<<q.nq, THR_THREADS, thr_smem, st>>>(S.samp, p.s_slices * (EPI_WARPS / 4), q.nq, p.nq_pad, p.r, r0, S.tau, S.qn, S.ctr + 1, metric, q.dpad,
                                                        p.stride, p.cap, (!we || atoi(we) != 0) ? 1 : 0);
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    const bool timing = stats && stats->time_scan && cudaEventCreate(&ev0) == cudaSuccess && cudaEventCreate(&ev1) == cudaSuccess;
    if (timing) cudaEventRecord(ev0, st);
    if (p.cluster) {
        CUtensorMap mapXh;
        if (!make_map(&mapXh, c.vecs, c.n, q.dpad, c.row_stride, BN / 2)) return cudaErrorNotSupported;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(p.nqt_launch, p.slices); cfg.blockDim = dim3(threads_of(true)); cfg.dynamicSmemBytes = SMEM_BYTES; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        TC_TRY(cudaLaunchKernelEx(&cfg, k_tc_scan<true, 1, true>, mapQ, mapXh, a));
    } else {
        k_tc_scan<true, 1><<<grid, threads_of(true), SMEM_BYTES, st>>>(mapQ, mapX, a);
    }
    if (timing) cudaEventRecord(ev1, st);
    if (a.flags & 24u) {   // rate measurement (bench.py: tcgen05 TF32 rate of this kernel without its epilogue / without its loads): nothing to re-rank
        TC_TRY(cudaStreamSynchronize(st));
        if (timing) { float ms = 0.f; if (cudaEventElapsedTime(&ms, ev0, ev1) == cudaSuccess) stats->scan_ms += ms; }
        if (stats) stats->scans++;
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        return cudaSuccess;
    }

    RerankArgs r{};
    r.vecs = c.vecs; r.sqn = c.sqn; r.slot_key = c.slot_key; r.deleted = c.deleted; r.first_slot = c.first_slot; r.n_slots = c.n_slots; r.dpad = q.dpad;
    r.q = q.q; r.q_stride = q.stride; r.qn = S.qn; r.nq = q.nq; r.qmap = nullptr; r.metric = metric; r.k = k; r.kp = p.kp; r.cap = p.cap;
    r.cand = S.cand; r.cand_d = S.cand_d; r.cnt = S.cnt; r.tau = S.tau; r.max_xn_bits = S.ctr + 1; r.out_keys = out_keys; r.out_slots = out_slots; r.out_dist = out_dist; r.out_ld = out_ld;
    r.self_first = q.self_first; r.self_step = q.self_step;
    r.n_r2 = S.ctr; r.r2_ids = S.r2_ids; r.r2_tau = S.r2_tau; r.r2_qn = S.r2_qn; r.n_fb = S.ctr + 2; r.fb_ids = S.fb_ids;
    r.n_active = nullptr; r.r2_cap = std::min<uint32_t>(q.nq, R2_MAX);
    k_tc_exact_dist<<<dim3(q.nq, (p.cap + 127) / 128), 128, dsmem, st>>>(r);
    auto select = [&](const RerankArgs& ra, uint32_t nqs) {
        if (c.slot_key) k_tc_select<LessByKey><<<(nqs + 3) / 4, 128, ssmem, st>>>(ra, LessByKey{c.slot_key, c.n_slots});
        else k_tc_select<LessById><<<(nqs + 3) / 4, 128, ssmem, st>>>(ra, LessById());
    };
    select(r, q.nq);
    TC_TRY(cudaGetLastError());
    if (stats) { stats->queries += q.nq; stats->scans++; }

    // Second round and fallback are launched UNCONDITIONALLY; how many queries they have to serve is read on the device
    // (S.ctr[0] = unproven queries with a proven threshold, S.ctr[2] = queries for the FP32 fallback). With nothing to do their
    // CTAs leave at once (~20 us for the lot), and the host never waits: vs_search_batch_device(exact = 1, stream) returns
    // before the work completes, and an insertion pipeline that asks for a batch's neighbours keeps enqueuing.
    {
        const uint32_t cap2 = std::min<uint32_t>(q.nq, R2_MAX);
        const Plan p2 = make_plan(cap2, c.n, kplan, q.dpad, dev.sm_count);
        k_tc_gather_queries<<<std::min<uint32_t>(cap2, 256u), 256, 0, st>>>(q.q, q.stride, q.dpad, S.r2_ids, S.ctr, cap2, S.gq);
        CUtensorMap mapQ2;
        if (!make_map(&mapQ2, S.gq, cap2, q.dpad, q.dpad, BM)) return cudaErrorNotSupported;
        TC_TRY(cudaMemsetAsync(S.cnt, 0, (size_t)cap2 * 4, st));
        ScanArgs a2 = a;
        a2.nq = cap2; a2.nq_pad = p2.nq_pad; a2.n_tiles = p2.n_tiles; a2.tiles_per_slice = p2.tiles_per_slice; a2.tau = S.r2_tau;
        a2.n_active = S.ctr; a2.active_cap = cap2;
        k_tc_scan<true, 1><<<dim3(p2.nqt, p2.slices), threads_of(true), SMEM_BYTES, st>>>(mapQ2, mapX, a2);
        RerankArgs r2 = r;
        r2.q = S.gq; r2.q_stride = q.dpad; r2.qn = S.r2_qn; r2.nq = cap2; r2.qmap = S.r2_ids; r2.tau = S.r2_tau; r2.n_r2 = nullptr; r2.n_active = S.ctr; r2.r2_cap = 0;
        k_tc_exact_dist<<<dim3(cap2, (p.cap + 127) / 128), 128, dsmem, st>>>(r2);
        select(r2, cap2);
        TC_TRY(cudaGetLastError());
        // what is left (fewer than k live candidates, overflow twice, more than R2_MAX unproven queries): the FP32 fallback
        const size_t fsmem = (size_t)q.dpad * 4 + 256 * 33 * 4 + 256 * 8 + (size_t)p.kp * 16 + 256;
        if (fsmem > dev.smem_per_block) return cudaErrorInvalidConfiguration;
        const uint32_t fblocks = std::min<uint32_t>(q.nq, (uint32_t)dev.sm_count * 2u);
        if (c.slot_key) {
            TC_TRY(cudaFuncSetAttribute(k_tc_fallback<LessByKey>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
            k_tc_fallback<LessByKey><<<fblocks, 256, fsmem, st>>>(r, c.n, LessByKey{c.slot_key, c.n_slots});
        } else {
            TC_TRY(cudaFuncSetAttribute(k_tc_fallback<LessById>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
            k_tc_fallback<LessById><<<fblocks, 256, fsmem, st>>>(r, c.n, LessById());
        }
        if (stats && stats->dev) k_tc_stats<<<1, 32, 0, st>>>(S.ctr, stats->dev);
        TC_TRY(cudaGetLastError());
    }
    (void)fb_scratch;
    if (timing) {   // profiling mode only (vs_set_stats): the scan kernel's own duration needs the events to have completed
        TC_TRY(cudaStreamSynchronize(st));
        float ms = 0.f; if (cudaEventElapsedTime(&ms, ev0, ev1) == cudaSuccess) stats->scan_ms += ms;
    }
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    return cudaSuccess;
#undef TC_TRY
}

}  // namespace vsann
