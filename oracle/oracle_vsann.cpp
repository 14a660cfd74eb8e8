// oracle_vsann.cpp — CPU restatement of the reference's hot path. TEST INFRASTRUCTURE ONLY.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; the product (vector-store-text_b200/csrc) never links, calls or falls back to it.
//
// PARITY UNPINNED at the arithmetic boundary: the reference delegates all ANN arithmetic to the
// third-party `usearch` crate, which is absent from the reference's Cargo.toml AND Cargo.lock
// (no version pinned anywhere) and from this machine; there is no Rust toolchain either
// (SURVEY.md F2/F3/F7, §8c). This file therefore restates
//   (i)  the glue semantics of src/index/usearch.rs:82-311 (key -> slot, replace = remove + add,
//        tombstone remove, count, reserve thresholds, dimension checks, result ordering), and
//   (ii) the published HNSW algorithm (Malkov & Yashunin 2018) as usearch v2.x applies it
//        (M / 2M link caps, level = floor(-ln U / ln M), greedy descent,
//        ef-wide beam, heuristic neighbour selection, reverse links with re-pruning, tombstones),
// and it is pinned against the only numeric expectations the reference holds for this path:
// the known-answer tests at src/index/usearch.rs:323-425 and tests/integration/usearch.rs:74-123
// (tests/test_oracle_kat.py), plus self-consistency with its own exact search.
//
// Exported ABI: include/vsann.h with an `o` prefix (ovs_*), so a parity test is
// "same calls, two backends".
//
// Distances (f32, the definition the GPU library must reproduce bit-for-bit in exact mode):
//   dot  = fmaf-chain over dimensions in index order, acc starts at 0
//   l2sq = fmaf(d, d, acc) with d = a[i] - b[i], in index order
//   cos  = 1 - dot / (sqrtf(na) * sqrtf(nb)), na/nb = fmaf-chain sums of squares;
//          both zero -> 0, one zero -> 1
//   ip   = 1 - dot
// Exact search orders by (distance asc, key asc).

#include <algorithm>
#include <atomic>
#include <cmath>
#if defined(__AVX2__) && defined(__FMA__)
#include <immintrin.h>
#define OVS_HAVE_SIMD 1
#else
#define OVS_HAVE_SIMD 0
#endif
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <mutex>
#include <queue>
#include <random>
#include <shared_mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../include/vsann.h"

namespace {

thread_local std::string g_err;
vs_status fail(vs_status s, const std::string& m) { g_err = m; return s; }

constexpr uint32_t NONE = 0xFFFFFFFFu;

inline float dot_seq(const float* a, const float* b, size_t d) {
    float acc = 0.f;
    for (size_t i = 0; i < d; ++i) acc = std::fmaf(a[i], b[i], acc);
    return acc;
}
inline float l2sq_seq(const float* a, const float* b, size_t d) {
    float acc = 0.f;
    for (size_t i = 0; i < d; ++i) { float t = a[i] - b[i]; acc = std::fmaf(t, t, acc); }
    return acc;
}
// 8-wide AVX2 + FMA distances with four independent accumulators: what a production CPU library (usearch links SimSIMD)
// runs per distance. OPT-IN (ovs_set_simd): only the timed CPU baseline of bench.py uses it, so that the baseline is not
// handicapped by the sequential chain above; every parity test keeps the sequential definition, and exact search always does.
#if OVS_HAVE_SIMD
inline float hsum8(__m256 v) {
    __m128 s = _mm_add_ps(_mm256_castps256_ps128(v), _mm256_extractf128_ps(v, 1));
    s = _mm_add_ps(s, _mm_movehl_ps(s, s));
    s = _mm_add_ss(s, _mm_shuffle_ps(s, s, 1));
    return _mm_cvtss_f32(s);
}
inline float dot_simd(const float* a, const float* b, size_t d) {
    __m256 s0 = _mm256_setzero_ps(), s1 = s0, s2 = s0, s3 = s0; size_t i = 0;
    for (; i + 32 <= d; i += 32) {
        s0 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i), s0);
        s1 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 8), _mm256_loadu_ps(b + i + 8), s1);
        s2 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 16), _mm256_loadu_ps(b + i + 16), s2);
        s3 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 24), _mm256_loadu_ps(b + i + 24), s3);
    }
    for (; i + 8 <= d; i += 8) s0 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i), s0);
    float acc = hsum8(_mm256_add_ps(_mm256_add_ps(s0, s1), _mm256_add_ps(s2, s3)));
    for (; i < d; ++i) acc = std::fmaf(a[i], b[i], acc);
    return acc;
}
inline float l2sq_simd(const float* a, const float* b, size_t d) {
    __m256 s0 = _mm256_setzero_ps(), s1 = s0, s2 = s0, s3 = s0; size_t i = 0;
    for (; i + 32 <= d; i += 32) {
        __m256 t0 = _mm256_sub_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i)), t1 = _mm256_sub_ps(_mm256_loadu_ps(a + i + 8), _mm256_loadu_ps(b + i + 8));
        __m256 t2 = _mm256_sub_ps(_mm256_loadu_ps(a + i + 16), _mm256_loadu_ps(b + i + 16)), t3 = _mm256_sub_ps(_mm256_loadu_ps(a + i + 24), _mm256_loadu_ps(b + i + 24));
        s0 = _mm256_fmadd_ps(t0, t0, s0); s1 = _mm256_fmadd_ps(t1, t1, s1); s2 = _mm256_fmadd_ps(t2, t2, s2); s3 = _mm256_fmadd_ps(t3, t3, s3);
    }
    for (; i + 8 <= d; i += 8) { __m256 t = _mm256_sub_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i)); s0 = _mm256_fmadd_ps(t, t, s0); }
    float acc = hsum8(_mm256_add_ps(_mm256_add_ps(s0, s1), _mm256_add_ps(s2, s3)));
    for (; i < d; ++i) { float t = a[i] - b[i]; acc = std::fmaf(t, t, acc); }
    return acc;
}
inline void prefetch_row(const float* v, size_t d) { for (size_t i = 0; i < d; i += 16) _mm_prefetch((const char*)(v + i), _MM_HINT_T0); }
#else
inline float dot_simd(const float* a, const float* b, size_t d) { return dot_seq(a, b, d); }
inline float l2sq_simd(const float* a, const float* b, size_t d) { return l2sq_seq(a, b, d); }
inline void prefetch_row(const float*, size_t) {}
#endif
inline float cos_from(float dot, float na, float nb) {
    if (na == 0.f && nb == 0.f) return 0.f;
    if (na == 0.f || nb == 0.f) return 1.f;
    return 1.f - dot / (std::sqrt(na) * std::sqrt(nb));
}

struct Node {
    int level = 0;
    std::vector<std::vector<uint32_t>> links;  // [level] -> slots
    std::mutex mu;
};

struct DistId {
    float d; uint32_t s;
};
struct FarFirst { bool operator()(const DistId& a, const DistId& b) const { return a.d < b.d || (a.d == b.d && a.s < b.s); } };
struct NearFirst { bool operator()(const DistId& a, const DistId& b) const { return a.d > b.d || (a.d == b.d && a.s > b.s); } };

}  // namespace

struct vs_index {
    size_t dims = 0; int metric = 0; size_t M = 16, efA = 128, efS = 64;
    size_t capacity = 0;
    std::vector<float> vecs;         // capacity x dims
    std::vector<float> sqn;          // squared norms (cos)
    std::vector<std::unique_ptr<Node>> nodes;
    std::vector<uint64_t> slot_key;
    std::vector<uint8_t> deleted;
    std::unordered_map<uint64_t, uint32_t> key_slot;
    std::atomic<uint32_t> n_slots{0};
    std::atomic<size_t> live{0};
    uint32_t entry = NONE; int max_level = -1;
    std::mutex entry_mu;
    std::shared_mutex big;           // exclusive: reserve / key map mutation; shared: add/search bodies
    std::mt19937_64 rng{0x5eed5eedULL};
    std::mutex rng_mu;
    double mL = 0;
    unsigned threads = 1;
    bool simd = false;               // ovs_set_simd: SIMD distances in the HNSW paths (timed CPU baseline only)
    std::atomic<uint64_t> c_sq{0}, c_sde{0}, c_sx{0}, c_bv{0}, c_bde{0};

    const float* vec(uint32_t s) const { return vecs.data() + (size_t)s * dims; }
    float dist_seq(const float* q, float qn, uint32_t s) const {   // the definition (exact search, every parity test)
        const float* v = vec(s);
        switch (metric) {
            case VS_METRIC_L2SQ: return l2sq_seq(q, v, dims);
            case VS_METRIC_IP: return 1.f - dot_seq(q, v, dims);
            default: return cos_from(dot_seq(q, v, dims), qn, sqn[s]);
        }
    }
    float dist(const float* q, float qn, uint32_t s) const {       // HNSW build / search
        if (!simd) return dist_seq(q, qn, s);
        const float* v = vec(s);
        switch (metric) {
            case VS_METRIC_L2SQ: return l2sq_simd(q, v, dims);
            case VS_METRIC_IP: return 1.f - dot_simd(q, v, dims);
            default: return cos_from(dot_simd(q, v, dims), qn, sqn[s]);
        }
    }
    float qnorm(const float* q) const { return metric == VS_METRIC_COS ? dot_seq(q, q, dims) : 0.f; }
    size_t max_links(int level) const { return level == 0 ? 2 * M : M; }
};

namespace {

struct Visited {
    std::vector<uint32_t> stamp; uint32_t cur = 0;
    void begin(size_t n) { if (stamp.size() < n) stamp.assign(n, 0), cur = 0; if (++cur == 0) { std::fill(stamp.begin(), stamp.end(), 0); cur = 1; } }
    bool test_set(uint32_t s) { if (stamp[s] == cur) return true; stamp[s] = cur; return false; }
};
thread_local Visited tl_vis;

// ef-wide beam on one level. `for_results`: tombstoned nodes are traversed but kept out of the
// result heap (Appendix B "traversed but never returned").
void search_layer(vs_index* ix, const float* q, float qn, std::vector<DistId>& eps, size_t ef, int level,
                  bool for_results, std::vector<DistId>& out, uint64_t& evals, uint64_t& expansions) {
    std::priority_queue<DistId, std::vector<DistId>, NearFirst> cand;
    std::priority_queue<DistId, std::vector<DistId>, FarFirst> top;
    Visited& vis = tl_vis; vis.begin(ix->capacity);
    for (auto& e : eps) {
        if (vis.test_set(e.s)) continue;
        cand.push(e);
        if (!(for_results && ix->deleted[e.s])) top.push(e);
    }
    std::vector<uint32_t> nb;
    while (!cand.empty()) {
        DistId c = cand.top();
        if (top.size() >= ef && c.d > top.top().d) break;
        cand.pop();
        ++expansions;
        Node& n = *ix->nodes[c.s];
        { std::lock_guard<std::mutex> g(n.mu); nb = n.links[level]; }
        if (ix->simd) for (uint32_t s : nb) if (vis.stamp[s] != vis.cur) prefetch_row(ix->vec(s), ix->dims);   // as usearch prefetches the neighbours of a node
        for (uint32_t s : nb) {
            if (vis.test_set(s)) continue;
            float d = ix->dist(q, qn, s); ++evals;
            if (top.size() < ef || d < top.top().d) {
                cand.push({d, s});
                if (!(for_results && ix->deleted[s])) { top.push({d, s}); if (top.size() > ef) top.pop(); }
            }
        }
    }
    out.resize(top.size());
    for (size_t i = top.size(); i-- > 0;) { out[i] = top.top(); top.pop(); }   // ascending
}

// HNSW heuristic: scan by ascending distance to the base; keep c iff it is closer to the base
// than to every already-kept neighbour.
void select_heuristic(vs_index* ix, const std::vector<DistId>& sorted, size_t m, std::vector<uint32_t>& keep, uint64_t& evals) {
    keep.clear();
    for (const DistId& c : sorted) {
        if (keep.size() >= m) break;
        bool good = true;
        const float* cv = ix->vec(c.s); float cn = ix->metric == VS_METRIC_COS ? ix->sqn[c.s] : 0.f;
        for (uint32_t k : keep) {
            float dk = ix->dist(cv, cn, k); ++evals;
            if (dk < c.d) { good = false; break; }
        }
        if (good) keep.push_back(c.s);
    }
}

int draw_level(vs_index* ix) {
    std::lock_guard<std::mutex> g(ix->rng_mu);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    double u = U(ix->rng); if (u <= 0) u = 1e-300;
    return (int)std::floor(-std::log(u) * ix->mL);
}

void hnsw_insert(vs_index* ix, uint32_t slot) {
    const float* q = ix->vec(slot); float qn = ix->metric == VS_METRIC_COS ? ix->sqn[slot] : 0.f;
    Node& me = *ix->nodes[slot];
    uint64_t evals = 0, exps = 0;
    std::unique_lock<std::mutex> eg(ix->entry_mu);
    int maxl = ix->max_level; uint32_t ep = ix->entry;
    if (ep == NONE) { ix->entry = slot; ix->max_level = me.level; return; }
    if (me.level <= maxl) eg.unlock();
    DistId cur{ix->dist(q, qn, ep), ep}; ++evals;
    for (int l = maxl; l > me.level; --l) {
        bool moved = true;
        while (moved) {
            moved = false;
            std::vector<uint32_t> nb;
            { Node& n = *ix->nodes[cur.s]; std::lock_guard<std::mutex> g(n.mu); nb = n.links[l]; }
            ++exps;
            for (uint32_t s : nb) { float d = ix->dist(q, qn, s); ++evals; if (d < cur.d) { cur = {d, s}; moved = true; } }
        }
    }
    std::vector<DistId> eps{cur}, W; std::vector<uint32_t> keep, keep2;
    for (int l = std::min(me.level, maxl); l >= 0; --l) {
        search_layer(ix, q, qn, eps, ix->efA, l, false, W, evals, exps);
        select_heuristic(ix, W, ix->M, keep, evals);
        { std::lock_guard<std::mutex> g(me.mu); me.links[l] = keep; }
        for (uint32_t nbs : keep) {
            Node& n = *ix->nodes[nbs]; std::lock_guard<std::mutex> g(n.mu);
            auto& L = n.links[l];
            if (L.size() < ix->max_links(l)) { L.push_back(slot); continue; }
            const float* nv = ix->vec(nbs); float nn = ix->metric == VS_METRIC_COS ? ix->sqn[nbs] : 0.f;
            std::vector<DistId> c; c.reserve(L.size() + 1);
            c.push_back({ix->dist(nv, nn, slot), slot}); ++evals;
            for (uint32_t s : L) { c.push_back({ix->dist(nv, nn, s), s}); ++evals; }
            std::sort(c.begin(), c.end(), [](const DistId& a, const DistId& b) { return a.d < b.d || (a.d == b.d && a.s < b.s); });
            select_heuristic(ix, c, ix->max_links(l), keep2, evals);
            L = keep2;
        }
        eps = W;
    }
    if (me.level > maxl) { ix->entry = slot; ix->max_level = me.level; }
    ix->c_bde += evals;
}

void hnsw_search(vs_index* ix, const float* q, size_t k, size_t ef, std::vector<DistId>& out) {
    uint64_t evals = 0, exps = 0; out.clear();
    uint32_t ep; int maxl;
    { std::lock_guard<std::mutex> g(ix->entry_mu); ep = ix->entry; maxl = ix->max_level; }
    if (ep == NONE) return;
    float qn = ix->qnorm(q);
    DistId cur{ix->dist(q, qn, ep), ep}; ++evals;
    for (int l = maxl; l > 0; --l) {
        bool moved = true;
        while (moved) {
            moved = false; std::vector<uint32_t> nb;
            { Node& n = *ix->nodes[cur.s]; std::lock_guard<std::mutex> g(n.mu); nb = n.links[l]; }
            ++exps;
            for (uint32_t s : nb) { float d = ix->dist(q, qn, s); ++evals; if (d < cur.d) { cur = {d, s}; moved = true; } }
        }
    }
    std::vector<DistId> eps{cur};
    search_layer(ix, q, qn, eps, std::max(ef, k), 0, true, out, evals, exps);
    if (out.size() > k) out.resize(k);
    ix->c_sde += evals; ix->c_sx += exps; ix->c_sq += 1;
}

void exact_search(vs_index* ix, const float* q, size_t k, std::vector<std::pair<float, uint64_t>>& out) {
    float qn = ix->qnorm(q); uint32_t n = ix->n_slots.load();
    out.clear(); out.reserve(ix->live.load());
    for (uint32_t s = 0; s < n; ++s) if (!ix->deleted[s]) out.push_back({ix->dist_seq(q, qn, s), ix->slot_key[s]});
    size_t kk = std::min(k, out.size());
    std::partial_sort(out.begin(), out.begin() + kk, out.end());
    out.resize(kk);
}

template <class F> void parallel_for(size_t n, unsigned threads, F f) {
    if (threads <= 1 || n < 2) { for (size_t i = 0; i < n; ++i) f(i); return; }
    std::atomic<size_t> next{0}; std::vector<std::thread> th;
    for (unsigned t = 0; t < threads; ++t) th.emplace_back([&] { for (;;) { size_t i = next.fetch_add(1); if (i >= n) break; f(i); } });
    for (auto& t : th) t.join();
}

vs_status do_reserve(vs_index* ix, size_t cap) {
    if (cap <= ix->capacity) return VS_OK;
    try {
        ix->vecs.resize(cap * ix->dims); ix->sqn.resize(cap); ix->nodes.resize(cap);
        ix->slot_key.resize(cap); ix->deleted.resize(cap, 0);
    } catch (...) { return fail(VS_ERR_CAPACITY, "reserve: out of memory"); }
    ix->capacity = cap; return VS_OK;
}

}  // namespace

extern "C" {

vs_status ovs_index_create(const vs_options* o, vs_index** out) {
    if (!o || !out) return fail(VS_ERR_INVALID, "create: null argument");
    if (o->dims == 0) return fail(VS_ERR_DIMS, "create: dimensions == 0");
    if (o->metric < 0 || o->metric > 2) return fail(VS_ERR_INVALID, "create: unknown metric");
    auto* ix = new vs_index();
    ix->dims = o->dims; ix->metric = o->metric;
    ix->M = o->connectivity ? o->connectivity : 16;
    ix->efA = o->expansion_add ? o->expansion_add : 128;
    ix->efS = o->expansion_search ? o->expansion_search : 64;
    if (ix->M < 2) ix->M = 2;
    ix->mL = 1.0 / std::log((double)ix->M);
    if (const char* e = getenv("OVS_SIMD")) ix->simd = atoi(e) != 0 && OVS_HAVE_SIMD;   // for callers that only hold the function table (lib/actor_load)
    vs_status s = do_reserve(ix, o->capacity_hint ? o->capacity_hint : 1000000);  // usearch.rs:62,99
    if (s != VS_OK) { delete ix; return s; }
    *out = ix; return VS_OK;
}
void ovs_index_destroy(vs_index* ix) { delete ix; }

vs_status ovs_set_threads(vs_index* ix, unsigned t) { if (!ix) return VS_ERR_INVALID; ix->threads = t ? t : 1; return VS_OK; }
vs_status ovs_set_simd(vs_index* ix, int on) { if (!ix) return VS_ERR_INVALID; ix->simd = on != 0 && OVS_HAVE_SIMD; return VS_OK; }
int ovs_have_simd(void) { return OVS_HAVE_SIMD; }
vs_status ovs_set_seed(vs_index* ix, uint64_t seed) { if (!ix) return VS_ERR_INVALID; ix->rng.seed(seed); return VS_OK; }

vs_status ovs_reserve(vs_index* ix, size_t cap) {
    if (!ix) return fail(VS_ERR_INVALID, "reserve: null index");
    std::unique_lock<std::shared_mutex> g(ix->big); return do_reserve(ix, cap);
}

vs_status ovs_remove(vs_index* ix, uint64_t key) {
    if (!ix) return fail(VS_ERR_INVALID, "remove: null index");
    std::unique_lock<std::shared_mutex> g(ix->big);
    auto it = ix->key_slot.find(key);
    if (it == ix->key_slot.end()) return fail(VS_NOT_FOUND, "remove: key not found");
    ix->deleted[it->second] = 1; ix->key_slot.erase(it); ix->live--; return VS_OK;
}

vs_status ovs_add_batch(vs_index* ix, size_t n, const uint64_t* keys, const float* vecs) {
    if (!ix || (n && (!keys || !vecs))) return fail(VS_ERR_INVALID, "add: null argument");
    std::unique_lock<std::shared_mutex> g(ix->big);
    // capacity growth rule of usearch.rs:201-212 (threshold 333 333, increment 1 000 000)
    while (ix->capacity - ix->n_slots.load() < std::max<size_t>(333333, n)) {
        vs_status s = do_reserve(ix, ix->capacity + std::max<size_t>(1000000, n)); if (s != VS_OK) return s;
    }
    uint32_t base = ix->n_slots.load();
    for (size_t i = 0; i < n; ++i) {
        uint32_t s = base + (uint32_t)i;
        auto it = ix->key_slot.find(keys[i]);
        if (it != ix->key_slot.end()) { ix->deleted[it->second] = 1; ix->live--; it->second = s; }   // replace = remove + add (usearch.rs:214-225)
        else ix->key_slot.emplace(keys[i], s);
        std::memcpy(ix->vecs.data() + (size_t)s * ix->dims, vecs + i * ix->dims, ix->dims * sizeof(float));
        ix->sqn[s] = dot_seq(ix->vec(s), ix->vec(s), ix->dims);
        ix->slot_key[s] = keys[i]; ix->deleted[s] = 0;
        auto nd = std::make_unique<Node>(); nd->level = draw_level(ix); nd->links.resize(nd->level + 1);
        ix->nodes[s] = std::move(nd);
    }
    ix->n_slots = base + (uint32_t)n; ix->live += n; ix->c_bv += n;
    size_t start = 0;
    if (ix->entry == NONE && n) { hnsw_insert(ix, base); start = 1; }
    parallel_for(n - start, ix->threads, [&](size_t i) { hnsw_insert(ix, base + (uint32_t)(start + i)); });
    return VS_OK;
}
vs_status ovs_add(vs_index* ix, uint64_t key, const float* v) { return ovs_add_batch(ix, 1, &key, v); }

vs_status ovs_search_batch(vs_index* ix, size_t nq, const float* qs, size_t k, size_t ef, int exact,
                           uint64_t* keys_out, float* dist_out, size_t* found) {
    if (!ix || !qs || !keys_out || !dist_out) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    std::shared_lock<std::shared_mutex> g(ix->big);
    size_t efe = ef ? ef : ix->efS;
    parallel_for(nq, ix->threads, [&](size_t i) {
        const float* q = qs + i * ix->dims; size_t f = 0;
        if (exact) {
            std::vector<std::pair<float, uint64_t>> r; exact_search(ix, q, k, r);
            for (; f < r.size(); ++f) { keys_out[i * k + f] = r[f].second; dist_out[i * k + f] = r[f].first; }
        } else {
            std::vector<DistId> r; hnsw_search(ix, q, k, efe, r);
            for (; f < r.size(); ++f) { keys_out[i * k + f] = ix->slot_key[r[f].s]; dist_out[i * k + f] = r[f].d; }
        }
        if (found) found[i] = f;
        for (; f < k; ++f) { keys_out[i * k + f] = UINT64_MAX; dist_out[i * k + f] = std::numeric_limits<float>::infinity(); }
    });
    return VS_OK;
}
vs_status ovs_search(vs_index* ix, const float* q, size_t k, uint64_t* ko, float* d, size_t* found) {
    return ovs_search_batch(ix, 1, q, k, 0, 0, ko, d, found);
}

size_t ovs_count(vs_index* ix) { return ix ? ix->live.load() : 0; }
size_t ovs_capacity(vs_index* ix) { return ix ? ix->capacity : 0; }
vs_status ovs_flush(vs_index*) { return VS_OK; }

vs_status ovs_get_stats(vs_index* ix, vs_stats* o) {
    if (!ix || !o) return VS_ERR_INVALID;
    std::memset(o, 0, sizeof(*o));
    o->count = ix->live; o->slots = ix->n_slots; o->capacity = ix->capacity;
    o->search_queries = ix->c_sq; o->search_dist_evals = ix->c_sde; o->search_expansions = ix->c_sx;
    o->build_vectors = ix->c_bv; o->build_dist_evals = ix->c_bde; o->top_level = ix->max_level < 0 ? 0 : ix->max_level; o->dims = (uint32_t)ix->dims;
    return VS_OK;
}
vs_status ovs_set_stats(vs_index*, int) { return VS_OK; }
vs_status ovs_reset_stats(vs_index* ix) { if (!ix) return VS_ERR_INVALID; ix->c_sq = ix->c_sde = ix->c_sx = ix->c_bv = ix->c_bde = 0; return VS_OK; }

// Stand-alone distance (tests pin Appendix C's table with it). metric as vs_metric.
float ovs_distance(int metric, const float* a, const float* b, size_t d) {
    switch (metric) {
        case VS_METRIC_L2SQ: return l2sq_seq(a, b, d);
        case VS_METRIC_IP: return 1.f - dot_seq(a, b, d);
        default: return cos_from(dot_seq(a, b, d), dot_seq(a, a, d), dot_seq(b, b, d));
    }
}

// Beam search on an EXTERNAL flat level-0 table (the GPU library's export): checks the CUDA
// traversal against the same algorithm on the same graph. links: n x width u32 (UINT32_MAX = empty),
// vecs: n x d, entry: start slot. Returns ascending (dist, slot) pairs; counts work.
vs_status ovs_graph_search(int metric, size_t d, size_t n, const float* vecs, const uint32_t* links, size_t width,
                           const uint8_t* deleted, uint32_t entry, const float* q, size_t k, size_t ef,
                           uint32_t* slots_out, float* dist_out, size_t* found, uint64_t* evals_out, uint64_t* exps_out) {
    if (!vecs || !links || !q || entry >= n) return fail(VS_ERR_INVALID, "graph_search: bad argument");
    auto dist = [&](uint32_t s) { return ovs_distance(metric, q, vecs + (size_t)s * d, d); };
    std::priority_queue<DistId, std::vector<DistId>, NearFirst> cand;
    std::priority_queue<DistId, std::vector<DistId>, FarFirst> top;
    std::vector<uint8_t> vis(n, 0); uint64_t ev = 1, ex = 0;
    DistId e{dist(entry), entry}; vis[entry] = 1; cand.push(e); top.push(e);
    ef = std::max(ef, k);
    while (!cand.empty()) {
        DistId c = cand.top(); if (top.size() >= ef && c.d > top.top().d) break; cand.pop(); ++ex;
        for (size_t j = 0; j < width; ++j) {
            uint32_t s = links[(size_t)c.s * width + j]; if (s == NONE || s >= n) continue;
            if (vis[s]) continue;
            vis[s] = 1; float dd = dist(s); ++ev;
            if (top.size() < ef || dd < top.top().d) { cand.push({dd, s}); top.push({dd, s}); if (top.size() > ef) top.pop(); }
        }
    }
    std::vector<DistId> r(top.size()); for (size_t i = top.size(); i-- > 0;) { r[i] = top.top(); top.pop(); }
    size_t f = 0;
    for (auto& x : r) { if (deleted && deleted[x.s]) continue; if (f >= k) break; slots_out[f] = x.s; dist_out[f] = x.d; ++f; }
    if (found) *found = f;
    if (evals_out) *evals_out = ev;
    if (exps_out) *exps_out = ex;
    return VS_OK;
}

const char* ovs_last_error(void) { return g_err.c_str(); }
const char* ovs_version(void) { return "vsann-oracle 0.1 cpu (test infrastructure)"; }

}  // extern "C"
