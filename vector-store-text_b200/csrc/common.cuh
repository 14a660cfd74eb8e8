// common.cuh — shared device primitives of the B200 ANN index (sm_100a).
//
// Layout in HBM (one index = one GPU shard):
//   vecs      f32  [capacity][dpad]       row-major, dpad = dims rounded up to 4 (zero padded), rows 16 B aligned
//   sqn       f32  [capacity]             squared norms, sequential fmaf chain (cosine)
//   links0    u32  [capacity][W0]         level-0 neighbour table, W0 = 2M, NONE-terminated rows (one 128 B line at M=16)
//   linksU[l] u32  [ceil(cap/M^l)][M]     level-l table; slot s lives on level l iff s % M^l == 0, row = s / M^l
//   slot_key  u64  [capacity]             slot -> caller key
//   deleted   u32  [capacity/32]          tombstone bitmask
//
// Levels are implicit in the slot number: P(level >= l) = M^-l, the HNSW distribution
// (reference semantics: SURVEY.md Appendix B), with no per-node level array and no pointer chase.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vsann {

constexpr uint32_t NONE = 0xFFFFFFFFu;
constexpr uint32_t EXPANDED = 0x80000000u;   // flag bit on list ids (slots < 2^31)
constexpr int MAX_LEVELS = 12;

enum Metric : int { COS = 0, L2SQ = 1, IP = 2 };

struct GraphView {
    const float* __restrict__ vecs;
    const float* __restrict__ sqn;
    const uint32_t* links0;
    const uint32_t* linksU[MAX_LEVELS];   // [0] unused
    uint32_t stride[MAX_LEVELS];          // M^l
    const uint32_t* __restrict__ deleted;
    const uint64_t* __restrict__ slot_key;
    uint32_t n_slots;                     // slots visible to this launch
    uint32_t dims, dpad, chunks;          // chunks = dpad / 4
    uint32_t M, W0;
    int metric;
    int top_level;
};

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ bool pair_less(float da, uint32_t ia, float db, uint32_t ib) {
    return da < db || (da == db && ia < ib);
}

__device__ __forceinline__ float4 ldg_f4(const float4* p) { return __ldg(p); }

// ---- metric finalisation -------------------------------------------------------------------
// `acc` is the dot product (COS/IP) or the squared distance (L2SQ).
__device__ __forceinline__ float finish_metric(int metric, float acc, float qn, float vn) {
    if (metric == L2SQ) return acc;
    if (metric == IP) return 1.f - acc;
    if (qn == 0.f && vn == 0.f) return 0.f;
    if (qn == 0.f || vn == 0.f) return 1.f;
    return 1.f - acc / (sqrtf(qn) * sqrtf(vn));
}

// One 16-byte chunk of a row against the query chunk, on Blackwell's packed-f32 pipe (FFMA2: two independent f32
// FMAs per instruction, each rounded like a scalar fmaf). The accumulator is a pair: .x sums the even components,
// .y the odd ones; the lane's partial sum is acc.x + acc.y. L2: t = q - x is formed as fma(x, -1, q) (exact product,
// one rounding: the same value as the subtraction).
template <int METRIC>
__device__ __forceinline__ void acc_chunk(const float4& q, const float4& x, float2& acc) {
    if (METRIC == L2SQ) {
        const float2 m1 = make_float2(-1.f, -1.f);
        const float2 t0 = __ffma2_rn(make_float2(x.x, x.y), m1, make_float2(q.x, q.y));
        const float2 t1 = __ffma2_rn(make_float2(x.z, x.w), m1, make_float2(q.z, q.w));
        acc = __ffma2_rn(t0, t0, acc);
        acc = __ffma2_rn(t1, t1, acc);
    } else {
        acc = __ffma2_rn(make_float2(q.x, q.y), make_float2(x.x, x.y), acc);
        acc = __ffma2_rn(make_float2(q.z, q.w), make_float2(x.z, x.w), acc);
    }
}

// ---- gathered distances --------------------------------------------------------------------
// One warp evaluates the distance from the query held in shared memory (q4, `chunks` float4s)
// to up to 32 gathered rows, one id per lane (NONE = skip). Rows are compacted, then teams of
// 8 lanes take one row each: lane t of a team loads float4 t, t+8, ... so every request of a
// team is one contiguous 128 B line; two rows per team are kept in flight.
// scr_id / scr_d: 32-entry per-warp shared scratch. Returns +inf for skipped lanes.
template <int METRIC>
__device__ __forceinline__ float warp_dist32_m(const GraphView& g, const float4* q4, float qn, uint32_t id,
                                               uint32_t* scr_id, float* scr_d, uint32_t& n_evals) {
    const unsigned lane = lane_id(), team = lane >> 3, tl = lane & 7u;
    const unsigned valid = __ballot_sync(0xffffffffu, id != NONE);
    const int nvalid = __popc(valid);
    const int pos = __popc(valid & ((1u << lane) - 1u));
    if (id != NONE) scr_id[pos] = id;
    __syncwarp();
    n_evals += nvalid;
    const int chunks = g.chunks;
    for (int base = 0; base < nvalid; base += 8) {
        const int j0 = base + team, j1 = base + 4 + team;
        const uint32_t a = j0 < nvalid ? scr_id[j0] : NONE;
        const uint32_t b = j1 < nvalid ? scr_id[j1] : NONE;
        const float4* va = reinterpret_cast<const float4*>(g.vecs + (size_t)(a == NONE ? 0 : a) * g.dpad);
        const float4* vb = reinterpret_cast<const float4*>(g.vecs + (size_t)(b == NONE ? 0 : b) * g.dpad);
        float2 pa = make_float2(0.f, 0.f), pb = make_float2(0.f, 0.f);
        if (a != NONE) {
#pragma unroll 4
            for (int c = tl; c < chunks; c += 8) {
                const float4 x = ldg_f4(va + c);
                float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
                if (b != NONE) y = ldg_f4(vb + c);
                const float4 q = q4[c];
                acc_chunk<METRIC>(q, x, pa); acc_chunk<METRIC>(q, y, pb);
            }
        }
        float acc_a = pa.x + pa.y, acc_b = pb.x + pb.y;
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) {
            acc_a += __shfl_xor_sync(0xffffffffu, acc_a, o);
            acc_b += __shfl_xor_sync(0xffffffffu, acc_b, o);
        }
        if (tl == 0) {
            if (a != NONE) scr_d[j0] = finish_metric(METRIC, acc_a, qn, METRIC == COS ? g.sqn[a] : 0.f);
            if (b != NONE) scr_d[j1] = finish_metric(METRIC, acc_b, qn, METRIC == COS ? g.sqn[b] : 0.f);
        }
    }
    __syncwarp();
    const float d = id != NONE ? scr_d[pos] : INFINITY;
    __syncwarp();
    return d;
}

__device__ __forceinline__ float warp_dist32(const GraphView& g, const float4* q4, float qn, uint32_t id,
                                             uint32_t* scr_id, float* scr_d, uint32_t& n_evals) {
    switch (g.metric) {
        case L2SQ: return warp_dist32_m<L2SQ>(g, q4, qn, id, scr_id, scr_d, n_evals);
        case IP: return warp_dist32_m<IP>(g, q4, qn, id, scr_id, scr_d, n_evals);
        default: return warp_dist32_m<COS>(g, q4, qn, id, scr_id, scr_d, n_evals);
    }
}

// Copy row `slot` into the per-warp shared query buffer; returns its squared norm (COS) or 0.
__device__ __forceinline__ float warp_load_row(const GraphView& g, float4* q4, uint32_t slot) {
    const float4* v = reinterpret_cast<const float4*>(g.vecs + (size_t)slot * g.dpad);
    for (int c = lane_id(); c < (int)g.chunks; c += 32) q4[c] = ldg_f4(v + c);
    __syncwarp();
    return g.metric == COS ? g.sqn[slot] : 0.f;
}

// ---- warp bitonic sort of 32 (dist, id) pairs, ascending, one per lane ----------------------
struct LessById {
    __device__ __forceinline__ bool operator()(float da, uint32_t ia, float db, uint32_t ib) const { return pair_less(da, ia, db, ib); }
};
// Exact search orders ties by caller key (oracle: distance asc, key asc).
struct LessByKey {
    const uint64_t* __restrict__ keys; uint32_t n;
    __device__ __forceinline__ bool operator()(float da, uint32_t ia, float db, uint32_t ib) const {
        if (da != db) return da < db;
        const uint64_t ka = ia < n ? keys[ia] : 0xFFFFFFFFFFFFFFFFull, kb = ib < n ? keys[ib] : 0xFFFFFFFFFFFFFFFFull;
        return ka < kb || (ka == kb && ia < ib);
    }
};

template <class Less>
__device__ __forceinline__ void warp_sort32(float& d, uint32_t& id, const Less& less) {
    const unsigned lane = lane_id();
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, d, j);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, id, j);
            const bool up = ((lane & k) == 0);          // ascending block
            const bool lower = ((lane & j) == 0);       // this lane keeps the smaller of the pair
            const bool take = (up == lower) ? less(od, oi, d, id) : less(d, id, od, oi);
            if (take) { d = od; id = oi; }
        }
    }
}

// ---- sorted candidate list in shared memory --------------------------------------------------
// The list holds the best `cap` (dist, id) pairs seen, ascending by (dist, id & ~EXPANDED);
// ld/li and ld2/li2 are ping-pong buffers of `capp` entries (capp = cap rounded up to 32; the
// tail stays (+inf, NONE)). merge_new folds <= 32 new candidates (one per lane, d = +inf /
// id = NONE to skip) and swaps the buffers. Duplicates of an id already in the list are dropped.
struct SList {
    float* ld; uint32_t* li; float* ld2; uint32_t* li2;
    int cap, capp;
};

__device__ __forceinline__ void slist_init(SList& L) {
    for (int i = lane_id(); i < L.capp; i += 32) { L.ld[i] = INFINITY; L.li[i] = NONE; L.ld2[i] = INFINITY; L.li2[i] = NONE; }
    __syncwarp();
}

__device__ __forceinline__ float slist_worst(const SList& L) { return L.ld[L.cap - 1]; }

template <class Less>
__device__ __forceinline__ void slist_merge(SList& L, float nd, uint32_t nid, uint32_t* scr_id, float* scr_d, const Less& less) {
    const unsigned lane = lane_id();
    if (nid == NONE) nd = INFINITY;
    if (__ballot_sync(0xffffffffu, nid != NONE) == 0u) return;
    warp_sort32(nd, nid, less);
    // position of each new candidate among the old entries (count of old < new), duplicate test
    int p = 0;
    bool ok = (nid != NONE);
    if (ok) {
        int lo = 0, hi = L.capp;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (less(L.ld[mid], L.li[mid] & ~EXPANDED, nd, nid)) lo = mid + 1; else hi = mid;
        }
        p = lo;
        if (p < L.capp && (L.li[p] & ~EXPANDED) == nid && L.li[p] != NONE) ok = false;   // already listed
    }
    const unsigned okm = __ballot_sync(0xffffffffu, ok);
    const int rank = __popc(okm & ((1u << lane) - 1u));
    const int nnew = __popc(okm);
    if (ok) { scr_d[rank] = nd; scr_id[rank] = nid; }
    __syncwarp();
    // old entries shift right by the number of new candidates that sort before them
    for (int i = lane; i < L.capp; i += 32) {
        const float od = L.ld[i]; const uint32_t oi = L.li[i];
        int lo = 0, hi = nnew;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (less(scr_d[mid], scr_id[mid], od, oi & ~EXPANDED)) lo = mid + 1; else hi = mid;
        }
        const int np = i + lo;
        if (np < L.cap) { L.ld2[np] = od; L.li2[np] = oi; }
    }
    if (ok) { const int np = p + rank; if (np < L.cap) { L.ld2[np] = nd; L.li2[np] = nid; } }
    __syncwarp();
    float* td = L.ld; L.ld = L.ld2; L.ld2 = td;
    uint32_t* ti = L.li; L.li = L.li2; L.li2 = ti;
}
__device__ __forceinline__ void slist_merge(SList& L, float nd, uint32_t nid, uint32_t* scr_id, float* scr_d) {
    slist_merge(L, nd, nid, scr_id, scr_d, LessById());
}

// First unexpanded entry among the best `cap`; marks it expanded. Returns NONE when the beam is done.
__device__ __forceinline__ uint32_t slist_pop(SList& L, float& d_out) {
    const unsigned lane = lane_id();
    for (int base = 0; base < L.cap; base += 32) {
        const int i = base + lane;
        const uint32_t id = i < L.cap ? L.li[i] : NONE;
        const bool open = (id != NONE) && !(id & EXPANDED);
        const unsigned m = __ballot_sync(0xffffffffu, open);
        if (m) {
            const int src = __ffs(m) - 1;
            const uint32_t pick = __shfl_sync(0xffffffffu, id, src);
            d_out = L.ld[base + src];
            if ((int)lane == src) L.li[i] = id | EXPANDED;
            __syncwarp();
            return pick;
        }
        // entries past the first NONE are all NONE
        if (__ballot_sync(0xffffffffu, i < L.cap && id == NONE)) break;
    }
    return NONE;
}

// ---- visited set: open-addressing hash in shared memory, lossy under pressure ---------------
// Returns true if `id` was already present. A full probe window evicts (forgets) an older id:
// the only cost is a redundant distance evaluation, duplicates are dropped by slist_merge.
__device__ __forceinline__ bool visited_test_set(uint32_t* hs, uint32_t hmask, uint32_t id) {
    uint32_t h = (id * 0x9E3779B1u) >> 7;
#pragma unroll 1
    for (int probe = 0; probe < 8; ++probe) {
        const uint32_t slot = (h + probe) & hmask;
        const uint32_t old = atomicCAS(&hs[slot], NONE, id);
        if (old == NONE) return false;
        if (old == id) return true;
    }
    hs[h & hmask] = id;
    return false;
}

// The same set with 16-bit TAGS, two per word (tables above 1024 slots: wide beams, where the table is most of a warp's shared
// memory - at ef = 384 it was 16 of 20 KB and held residency to 8 warps per SM). Slot = hash(id), tag = 15 other hash bits | 1
// (0 = empty). A foreign id with the same tag in the probe window reads as "visited" (p < 8 / 32768 per lookup: one candidate in
// ~4000 is skipped); everything else behaves like the 32-bit table, evictions included.
__device__ __forceinline__ bool visited_test_set16(uint32_t* hw, uint32_t smask, uint32_t id) {
    const uint32_t h = (id * 0x9E3779B1u) >> 7, tag = ((id * 0x85EBCA6Bu) >> 16) | 1u;
#pragma unroll 1
    for (int probe = 0; probe < 8; ++probe) {
        const uint32_t s = (h + probe) & smask, sh = (s & 1u) * 16u;
        uint32_t* w = hw + (s >> 1);
        uint32_t cur = *reinterpret_cast<volatile uint32_t*>(w);
        for (;;) {
            const uint32_t t = (cur >> sh) & 0xFFFFu;
            if (t == tag) return true;
            if (t != 0u) break;                                   // somebody else's slot: next probe
            const uint32_t prev = atomicCAS(w, cur, cur | (tag << sh));
            if (prev == cur) return false;
            cur = prev;                                           // the other half of the word changed meanwhile: look again
        }
    }
    const uint32_t s = h & smask, sh = (s & 1u) * 16u;            // full window: forget the entry of the home slot
    uint32_t* w = hw + (s >> 1);
    const uint32_t cur = *reinterpret_cast<volatile uint32_t*>(w);
    atomicCAS(w, cur, (cur & ~(0xFFFFu << sh)) | (tag << sh));    // best effort: losing this race only costs a re-evaluation later
    return false;
}
// Which form a launch uses travels in the top bit of the `hsize` kernel argument (HASH_TAGS | slots), so that the host's
// shared-memory arithmetic and the device's carving cannot disagree. The launcher sets it when the 32-bit table would be at
// least as large as the warp's staging area (hash_size_for, graph_kernels.cu): measured on B200, tags buy residency there
// (12.5M x 96: ef 384-1024 x1.4-1.6, ef 192-256 +5 %; 5M x 1536 ef 768-1024 x1.7) and cost 8 % where the staging area dominates
// anyway (5M x 1536, ef 192-512: a tag probe is a load + a CAS where the 32-bit probe is one CAS).
constexpr uint32_t HASH_TAGS = 0x80000000u;
__host__ __device__ inline bool hash_uses_tags(uint32_t hsize) { return (hsize & HASH_TAGS) != 0u; }
__host__ __device__ inline uint32_t hash_slots(uint32_t hsize) { return hsize & ~HASH_TAGS; }

__device__ __forceinline__ const uint32_t* level_row(const GraphView& g, int level, uint32_t slot, int& width) {
    if (level == 0) { width = g.W0; return g.links0 + (size_t)slot * g.W0; }
    width = g.M;
    return g.linksU[level] + (size_t)(slot / g.stride[level]) * g.M;
}

__device__ __forceinline__ bool is_deleted(const GraphView& g, uint32_t slot) {
    return (g.deleted[slot >> 5] >> (slot & 31u)) & 1u;
}

}  // namespace vsann
