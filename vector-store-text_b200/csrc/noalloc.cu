// noalloc.cu — vs_exact_search_noalloc_device: brute-force k-NN that allocates NOTHING (include/vsann.h).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). It is a variant of the FP32 exact search
// K1f (exact_kernels.cu: k_exact_tiles + k_exact_final), which keeps per-chunk partial top-k lists in scratch memory: here
// that scratch is the file-scope array below, VS_NOALLOC_SCRATCH_BYTES = 8 MiB, held by every GPU that loads the library
// until the process exits. Same arithmetic as the oracle's exact search (oracle/oracle_vsann.cpp exact_search): one
// sequential fmaf chain per (query, row) over the dimensions in index order, results ordered by (distance, key) - so the
// output carries the same bits as vs_search_batch(exact = 1).
//
// The array's NAME is a host-side stand-in in host code, not a device address: the address every launch gets comes from
// cudaGetSymbolAddress, asked once per device with that device current (noalloc_scratch()).
// Every call shares the array and finds the previous call's contents in it. Nothing in it has to start at zero: kernel A
// writes all kp entries of every (query, chunk) list - padding included - before kernel B reads them, in stream order, so a
// call never reads a word it has not written itself. Calls that OVERLAP on the GPU (two streams, two host threads) race on
// the array: the caller serialises them.
#include <mutex>

#include "shard.cuh"

namespace vsann {

constexpr uint32_t NA_CHUNKS = 64;                       // corpus chunks per query = partial lists per query
constexpr uint32_t NA_THREADS = 256;
constexpr uint32_t NA_KP_MAX = VS_NOALLOC_MAX_K;         // 256
constexpr size_t NA_SCRATCH_BYTES = VS_NOALLOC_SCRATCH_BYTES;
static_assert((size_t)VS_NOALLOC_MAX_QUERY_K * NA_CHUNKS * 8 == NA_SCRATCH_BYTES, "limits and scratch size disagree");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_SCRATCH_BYTES];

struct NaArgs {
    const float* vecs; const float* queries; const uint64_t* keys;   // keys may be null: the key of row r is r
    uint32_t n, dims, nq, k, kp, rows_per_chunk; int metric;
    float* part_d; uint32_t* part_i;                                 // [nq][NA_CHUNKS][kp], inside the static scratch
    uint64_t* keys_out; float* dist_out;
};

// Kernel A: block (chunk, query). Thread-per-row sequential chains; rows that can still enter the block's top-k are compacted
// into a 256-entry buffer and merged into the sorted list (common.cuh SList) by warp 0. Writes the whole kp-entry list.
template <class Less>
__global__ void __launch_bounds__(NA_THREADS) k_noalloc_partial(NaArgs a, Less less) {
    extern __shared__ __align__(16) float na_q[];                    // the query, dims floats
    __shared__ float bd[NA_THREADS]; __shared__ uint32_t bi[NA_THREADS];
    __shared__ float l_d[2][NA_KP_MAX]; __shared__ uint32_t l_i[2][NA_KP_MAX];
    __shared__ uint32_t scr_id[32]; __shared__ float scr_d[32];
    __shared__ uint32_t nbuf; __shared__ float worst;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u, chunk = blockIdx.x, q = blockIdx.y;
    for (uint32_t c = tid; c < a.dims; c += NA_THREADS) na_q[c] = a.queries[(size_t)q * a.dims + c];
    SList L; L.ld = l_d[0]; L.ld2 = l_d[1]; L.li = l_i[0]; L.li2 = l_i[1]; L.cap = (int)a.k; L.capp = (int)a.kp;
    if (warp == 0) slist_init(L);
    if (tid == 0) { nbuf = 0; worst = INFINITY; }
    __syncthreads();
    float qn = 0.f;
    if (a.metric == COS) for (uint32_t c = 0; c < a.dims; ++c) qn = fmaf(na_q[c], na_q[c], qn);   // the oracle's dot_seq(q, q)
    const uint32_t r_begin = min(a.n, chunk * a.rows_per_chunk), r_end = min(a.n, r_begin + a.rows_per_chunk);
    for (uint32_t r0 = r_begin; r0 < r_end; r0 += NA_THREADS) {      // block-uniform trip count
        const uint32_t row = r0 + tid; const bool valid = row < r_end;
        float acc = 0.f, xn = 0.f;
        if (valid) {
            const float* x = a.vecs + (size_t)row * a.dims;
            if (a.metric == L2SQ) { for (uint32_t c = 0; c < a.dims; ++c) { const float t = na_q[c] - __ldg(x + c); acc = fmaf(t, t, acc); } }
            else if (a.metric == IP) { for (uint32_t c = 0; c < a.dims; ++c) acc = fmaf(na_q[c], __ldg(x + c), acc); }
            else { for (uint32_t c = 0; c < a.dims; ++c) { const float v = __ldg(x + c); acc = fmaf(na_q[c], v, acc); xn = fmaf(v, v, xn); } }
        }
        const float d = valid ? finish_metric(a.metric, acc, qn, xn) : INFINITY;
        const bool ok = valid && d <= worst;                         // ties are admitted: the key order decides in the merge
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        uint32_t base = 0;
        if (lane == 0 && m) base = atomicAdd(&nbuf, (uint32_t)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (ok) { const uint32_t pos = base + __popc(m & ((1u << lane) - 1u)); bd[pos] = d; bi[pos] = row; }
        __syncthreads();
        if (warp == 0) {
            const uint32_t nb = nbuf;
            for (uint32_t b0 = 0; b0 < nb; b0 += 32) {
                const bool on = b0 + lane < nb;
                slist_merge(L, on ? bd[b0 + lane] : INFINITY, on ? bi[b0 + lane] : NONE, scr_id, scr_d, less);
            }
            if (lane == 0) { worst = L.ld[a.k - 1]; nbuf = 0; }      // slist_merge swaps L.ld / L.ld2: L lives in warp 0's registers
        }
        __syncthreads();
    }
    if (warp == 0) {                                                 // every entry, padding (+inf, NONE) included
        const size_t o = ((size_t)q * NA_CHUNKS + chunk) * a.kp;
        for (uint32_t i = lane; i < a.kp; i += 32) { a.part_d[o + i] = L.ld[i]; a.part_i[o + i] = L.li[i]; }
    }
}

// Kernel B: one warp per query folds its NA_CHUNKS partial lists into the final ascending top-k by (distance, key).
template <class Less>
__global__ void __launch_bounds__(128) k_noalloc_final(NaArgs a, Less less) {
    extern __shared__ __align__(16) unsigned char na_sm[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, q = blockIdx.x * (blockDim.x >> 5) + warp;
    if (q >= a.nq) return;
    unsigned char* base = na_sm + (size_t)warp * ((size_t)a.kp * 16 + 256);
    SList L;
    L.ld = reinterpret_cast<float*>(base); L.ld2 = L.ld + a.kp; L.li = reinterpret_cast<uint32_t*>(L.ld2 + a.kp); L.li2 = L.li + a.kp;
    uint32_t* scr_id = L.li2 + a.kp; float* scr_d = reinterpret_cast<float*>(scr_id + 32);
    L.cap = (int)a.k; L.capp = (int)a.kp;
    slist_init(L);
    const size_t o = (size_t)q * NA_CHUNKS * a.kp;
    const uint32_t total = NA_CHUNKS * a.kp;
    for (uint32_t b0 = 0; b0 < total; b0 += 32) {
        const uint32_t id = a.part_i[o + b0 + lane];
        const float d = id != NONE ? a.part_d[o + b0 + lane] : INFINITY;
        slist_merge(L, d, id, scr_id, scr_d, less);
    }
    for (uint32_t i = lane; i < a.k; i += 32) {
        const uint32_t id = L.li[i];
        a.keys_out[(size_t)q * a.k + i] = id == NONE ? 0xFFFFFFFFFFFFFFFFull : (a.keys ? a.keys[id] : (uint64_t)id);
        a.dist_out[(size_t)q * a.k + i] = id == NONE ? INFINITY : L.ld[i];
    }
}

// Device address of the static scratch on `device` (which must be current): asked once per device, then cached.
static vs_status noalloc_scratch(int device, unsigned char** out) {
    static std::mutex mu; static unsigned char* ptr[64] = {};
    if (device < 0 || device >= 64) return fail(VS_ERR_INVALID, "no-alloc search: bad device ordinal");
    std::lock_guard<std::mutex> g(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) { cudaGetLastError(); return fail(VS_ERR_CUDA, std::string("no-alloc search: cudaGetSymbolAddress: ") + cudaGetErrorString(e)); }
        ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = ptr[device];
    return VS_OK;
}

template <class Less>
static cudaError_t noalloc_launch(const NaArgs& a, Less less, cudaStream_t st) {
    k_noalloc_partial<Less><<<dim3(NA_CHUNKS, a.nq), NA_THREADS, (size_t)a.dims * 4, st>>>(a, less);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const uint32_t wpb = 4;
    k_noalloc_final<Less><<<(a.nq + wpb - 1) / wpb, wpb * 32, (size_t)wpb * ((size_t)a.kp * 16 + 256), st>>>(a, less);
    return cudaGetLastError();
}

}  // namespace vsann

using namespace vsann;

extern "C" vs_status vs_exact_search_noalloc_device(int device, int32_t metric, size_t dims, size_t n, const float* d_vecs, const uint64_t* d_keys,
                                                    size_t nq, const float* d_queries, size_t k, uint64_t* d_keys_out, float* d_dist_out, void* stream) {
    if (!d_vecs || !d_queries || !d_keys_out || !d_dist_out) return fail(VS_ERR_INVALID, "no-alloc search: null argument");
    if (metric < 0 || metric > 2) return fail(VS_ERR_INVALID, "no-alloc search: unknown metric");
    if (dims == 0 || dims > VS_NOALLOC_MAX_DIMS) return fail(VS_ERR_DIMS, "no-alloc search: dimensions must be in [1, " + std::to_string(VS_NOALLOC_MAX_DIMS) + "]");
    if (k == 0 || k > VS_NOALLOC_MAX_K) return fail(VS_ERR_UNSUPPORTED, "no-alloc search: k must be in [1, " + std::to_string(VS_NOALLOC_MAX_K) + "], got " + std::to_string(k));
    if (n == 0 || n > 0x7FFFFFF0ull) return fail(VS_ERR_UNSUPPORTED, "no-alloc search: rows must be in [1, 2^31 - 16]");
    if (nq == 0) return VS_OK;
    const size_t kp = (k + 31) & ~(size_t)31;
    if (nq > 65535 || nq * kp > VS_NOALLOC_MAX_QUERY_K)
        return fail(VS_ERR_UNSUPPORTED, "no-alloc search: queries x k (k rounded up to 32) = " + std::to_string(nq * kp) + " exceeds the static scratch limit of " +
                                        std::to_string(VS_NOALLOC_MAX_QUERY_K) + " (" + std::to_string(VS_NOALLOC_SCRATCH_BYTES >> 20) + " MiB): split the query block");
    CU(cudaSetDevice(device));
    unsigned char* scratch = nullptr;
    vs_status s = noalloc_scratch(device, &scratch); if (s != VS_OK) return s;
    NaArgs a{};
    a.vecs = d_vecs; a.queries = d_queries; a.keys = d_keys; a.n = (uint32_t)n; a.dims = (uint32_t)dims; a.nq = (uint32_t)nq; a.k = (uint32_t)k; a.kp = (uint32_t)kp;
    a.metric = metric;
    const uint32_t per = (uint32_t)((n + NA_CHUNKS - 1) / NA_CHUNKS);
    a.rows_per_chunk = ((per + NA_THREADS - 1) / NA_THREADS) * NA_THREADS;
    a.part_d = reinterpret_cast<float*>(scratch);
    a.part_i = reinterpret_cast<uint32_t*>(scratch + nq * NA_CHUNKS * kp * 4);      // <= 8 MiB in total by the check above
    a.keys_out = d_keys_out; a.dist_out = d_dist_out;
    cudaStream_t st = (cudaStream_t)stream;
    if (d_keys) CU(noalloc_launch(a, LessByKey{d_keys, (uint32_t)n}, st));
    else CU(noalloc_launch(a, LessById(), st));
    if (!stream) CU(cudaStreamSynchronize(nullptr));
    return VS_OK;
}
