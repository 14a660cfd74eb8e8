// index.cu — host side of libvsann.so: the C ABI of include/vsann.h over the sm_100a kernels.
//
// Mirrors what the reference's index actor does around usearch (src/index/usearch.rs):
//   new + reserve           usearch.rs:82-99      -> vs_index_create
//   add / replace           usearch.rs:174-233    -> vs_add*, staged and landed in batches
//   remove (tombstone)      usearch.rs:235-249    -> vs_remove
//   search -> Matches       usearch.rs:251-306    -> vs_search*
//   size / capacity/reserve usearch.rs:201-212,308-311 -> vs_count / vs_capacity / vs_reserve
// There is no CPU path in this library: every operation that touches vectors runs on the GPU.
#include <algorithm>
#include <condition_variable>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/vsann.h"
#include "kernels.h"

using namespace vsann;

namespace vsann {
__global__ void k_pad(const float* __restrict__ src, float* __restrict__ dst, uint32_t n, uint32_t dims, uint32_t dpad) {
    const size_t total = (size_t)n * dpad;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(i / dpad), c = (uint32_t)(i % dpad);
        dst[i] = c < dims ? src[(size_t)r * dims + c] : 0.f;
    }
}
}
static cudaError_t pad_queries(const float* src, float* dst, uint32_t n, uint32_t dims, uint32_t dpad, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const size_t total = (size_t)n * dpad;
    uint32_t blocks = (uint32_t)std::min<size_t>((total + 255) / 256, 148 * 8);
    vsann::k_pad<<<blocks, 256, 0, st>>>(src, dst, n, dims, dpad);
    return cudaGetLastError();
}

namespace {
thread_local std::string g_err;
vs_status fail(vs_status s, const std::string& m) { g_err = m; return s; }
#define CU(expr)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) return fail(VS_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)

template <class T> struct DBuf {
    T* p = nullptr; size_t n = 0;
    cudaError_t ensure(size_t want) {
        if (want <= n) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; n = 0;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) n = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};
}  // namespace

struct vs_index {
    // options
    uint32_t dims = 0, dpad = 0; int metric = 0; uint32_t M = 16, W0 = 32, efA = 128, efS = 64;
    uint32_t max_batch = 32768, batch_div = 16; size_t reserve_increment = 1000000;
    DeviceInfo dev; cudaStream_t stream = nullptr, side = nullptr; cudaEvent_t ev0 = nullptr, ev1 = nullptr, user_ev = nullptr, fork_ev = nullptr, join_ev = nullptr; bool user_pending = false, build_overlap = true;
    std::mutex mu;
    // tables
    size_t capacity = 0; uint32_t n_slots = 0; size_t live = 0; int top_level = 0; int n_levels = 1;
    float* vecs = nullptr; float* sqn = nullptr; uint32_t* links0 = nullptr; uint32_t* linksU[MAX_LEVELS] = {};
    uint64_t* slot_key = nullptr; uint32_t* deleted = nullptr;
    uint32_t stride[MAX_LEVELS] = {}; size_t rows[MAX_LEVELS] = {};
    LinkTables lt{}; size_t total_rows = 0; size_t edge_cap = 0;
    // build scratch
    BuildBatch bb{}; uint32_t efpA = 128; DBuf<unsigned char> build_scratch, exact_scratch, tc_scratch, stage; TcStats tc_stats, tc_build_stats; bool build_tc = true, zero_copy = true; uint64_t st_stream_queries = 0;
    DBuf<uint32_t> tomb; DBuf<float> qpad, qdense, out_d; DBuf<uint64_t> out_k;
    unsigned long long* counters = nullptr; bool stats_on = false;
    // host state
    std::unordered_map<uint64_t, uint32_t> key_slot;
    std::vector<uint64_t> pend_keys; std::vector<float> pend_vecs; std::unordered_map<uint64_t, size_t> pend_pos;
    std::vector<uint64_t> host_slot_key;
    uint64_t st_build_vectors = 0, st_build_batches = 0; double st_build_seconds = 0; size_t device_bytes = 0;
    static constexpr int KNN_SETS = 3;   // neighbour buffers: batch i links from one while the side stream fills the next two
    uint32_t* knn_set_id[KNN_SETS][MAX_LEVELS] = {}; float* knn_set_d[KNN_SETS][MAX_LEVELS] = {}; cudaEvent_t knn_done[KNN_SETS] = {}, link_done[KNN_SETS] = {}; int build_pipeline = 1;   // batches of look-ahead for the intra-batch neighbours (2 measured equal to 1)
    bool build_trace = false; std::vector<cudaEvent_t> trace_ev;   // VSANN_BUILD_TRACE=1: four events per batch (start, search, batch-knn, link)
    // vs_search coalescer: single-query calls that arrive while a launch is in flight ride the next launch together
    struct SearchReq { const float* q; size_t k; uint64_t* keys; float* dist; size_t* found; vs_status st; std::string err; bool done; };
    std::mutex co_mu; std::condition_variable co_cv; std::vector<SearchReq*> co_q; bool co_leader = false, coalesce = true;
    DBuf<uint32_t> found_d; uint32_t* found_pin = nullptr; size_t found_cap = 0;   // per-query result counts (vs_search_batch's `found`)
    float* co_pin_q = nullptr; uint64_t* co_pin_k = nullptr; float* co_pin_d = nullptr; size_t co_cap_q = 0, co_cap_o = 0;
    uint64_t st_search_calls = 0, st_search_launches = 0;

    GraphView view(uint32_t visible, int top) const {
        GraphView g{};
        g.vecs = vecs; g.sqn = sqn; g.links0 = links0;
        for (int l = 0; l < MAX_LEVELS; ++l) { g.linksU[l] = linksU[l]; g.stride[l] = stride[l] ? stride[l] : 0xFFFFFFFFu; }
        g.deleted = deleted; g.slot_key = slot_key; g.n_slots = visible;
        g.dims = dims; g.dpad = dpad; g.chunks = dpad / 4; g.M = M; g.W0 = W0; g.metric = metric; g.top_level = top;
        return g;
    }
};

namespace {

int top_for(const vs_index* ix, uint32_t n) {
    int t = 0;
    for (int l = 1; l < ix->n_levels; ++l) if (n > 0 && ix->stride[l] <= n - 1) t = l;
    return t;
}

template <class T> cudaError_t grow(T*& p, size_t old_n, size_t new_n, int fill_byte, cudaStream_t st, size_t& bytes) {
    T* np = nullptr;
    cudaError_t e = cudaMalloc(&np, new_n * sizeof(T));
    if (e != cudaSuccess) return e;
    if (fill_byte >= 0) { e = cudaMemsetAsync(np, fill_byte, new_n * sizeof(T), st); if (e != cudaSuccess) { cudaFree(np); return e; } }
    if (p && old_n) { e = cudaMemcpyAsync(np, p, std::min(old_n, new_n) * sizeof(T), cudaMemcpyDeviceToDevice, st); if (e != cudaSuccess) { cudaFree(np); return e; } }
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { cudaFree(np); return e; }
    if (p) cudaFree(p);
    bytes += (new_n - (p ? old_n : 0)) * sizeof(T);
    p = np;
    return cudaSuccess;
}

// usearch::Index::reserve (usearch.rs:99,206): grows every table; contents are preserved.
vs_status do_reserve(vs_index* ix, size_t cap) {
    if (cap <= ix->capacity) return VS_OK;
    if (cap >= 0x7FFFFFF0ull) return fail(VS_ERR_CAPACITY, "reserve: capacity exceeds 2^31 slots per shard");
    cudaSetDevice(ix->dev.device);
    const size_t old = ix->capacity;
    cudaError_t e;
#define GROW(call) do { e = (call); if (e != cudaSuccess) { cudaGetLastError(); return fail(VS_ERR_CAPACITY, std::string("reserve: ") + cudaGetErrorString(e)); } } while (0)
    GROW(grow(ix->vecs, old * ix->dpad, cap * ix->dpad, -1, ix->stream, ix->device_bytes));
    GROW(grow(ix->sqn, old, cap, -1, ix->stream, ix->device_bytes));
    GROW(grow(ix->links0, old * ix->W0, cap * ix->W0, 0xFF, ix->stream, ix->device_bytes));
    GROW(grow(ix->slot_key, old, cap, 0xFF, ix->stream, ix->device_bytes));
    GROW(grow(ix->deleted, (old + 31) / 32, (cap + 31) / 32, 0, ix->stream, ix->device_bytes));
    // level tables: level l exists while M^l < capacity
    int nl = 1; uint64_t s = ix->M;
    while (nl < MAX_LEVELS && s < cap) { ix->stride[nl] = (uint32_t)s; ++nl; s *= ix->M; }
    ix->stride[0] = 1;
    size_t total = cap; ix->lt.row_base[0] = 0;
    for (int l = 1; l < nl; ++l) {
        const size_t nr = (cap + ix->stride[l] - 1) / ix->stride[l];
        GROW(grow(ix->linksU[l], ix->rows[l] * ix->M, nr * ix->M, 0xFF, ix->stream, ix->device_bytes));
        ix->rows[l] = nr; ix->lt.row_base[l] = (uint32_t)total; total += nr;
    }
    ix->rows[0] = cap; ix->n_levels = nl;
    // inbox index space covers all rows; it is all-zero between batches so no copy is needed
    if (ix->lt.inbox_cnt) { cudaFree(ix->lt.inbox_cnt); cudaFree(ix->lt.inbox_idx); ix->device_bytes -= ix->total_rows * 8; ix->lt.inbox_cnt = ix->lt.inbox_idx = nullptr; }
    GROW(cudaMalloc(&ix->lt.inbox_cnt, total * 4)); GROW(cudaMalloc(&ix->lt.inbox_idx, total * 4));
    GROW(cudaMemsetAsync(ix->lt.inbox_cnt, 0, total * 4, ix->stream)); GROW(cudaMemsetAsync(ix->lt.inbox_idx, 0xFF, total * 4, ix->stream));
    GROW(cudaStreamSynchronize(ix->stream));
    ix->total_rows = total; ix->device_bytes += total * 8;
#undef GROW
    ix->capacity = cap;
    ix->host_slot_key.resize(cap);
    return VS_OK;
}

// Carves the per-batch scratch (candidate lists, intra-batch neighbours, edge/inbox arrays).
vs_status setup_build_scratch(vs_index* ix) {
    const uint32_t B = ix->max_batch, efp = build_efp(ix->efA), kbp = KNN_BATCH;
    ix->efpA = efp;
    size_t bytes = 0; size_t off_ci[MAX_LEVELS], off_cd[MAX_LEVELS], off_cn[MAX_LEVELS], off_ki[MAX_LEVELS], off_kd[MAX_LEVELS], off_kis[2][MAX_LEVELS], off_kds[2][MAX_LEVELS];
    auto take = [&](size_t b) { size_t o = bytes; bytes += (b + 255) & ~(size_t)255; return o; };
    size_t items_total = 0; uint64_t s = 1;
    for (int l = 0; l < MAX_LEVELS; ++l) {
        const size_t it = B / s + 2; items_total += it;
        off_ci[l] = take(it * efp * 4); off_cd[l] = take(it * efp * 4); off_cn[l] = take(it * 4);
        off_ki[l] = take(it * kbp * 4); off_kd[l] = take(it * kbp * 4);
        for (int e = 0; e < 2; ++e) { off_kis[e][l] = take(it * kbp * 4); off_kds[e][l] = take(it * kbp * 4); }   // further sets: the neighbours of batches i+1, i+2 are computed while batch i searches and links
        s *= ix->M; if (s > 0xFFFFFFFFull) s = 0xFFFFFFFFull;
    }
    ix->edge_cap = items_total * ix->M;
    const size_t ec = ix->edge_cap;
    const size_t o_ne = take(8), o_er = take(ec * 4), o_es = take(ec * 4), o_tr = take(ec * 4), o_ts = take(ec * 4), o_tl = take(ec),
                 o_if = take(ec * 4), o_in = take(ec * INBOX_CAP * 4);
    CU(ix->build_scratch.ensure(bytes));
    unsigned char* p = ix->build_scratch.p;
    CU(cudaMemsetAsync(p, 0, bytes, ix->stream));
    for (int l = 0; l < MAX_LEVELS; ++l) {
        ix->bb.cand_id[l] = (uint32_t*)(p + off_ci[l]); ix->bb.cand_d[l] = (float*)(p + off_cd[l]); ix->bb.cand_n[l] = (uint32_t*)(p + off_cn[l]);
        ix->bb.knn_id[l] = (uint32_t*)(p + off_ki[l]); ix->bb.knn_d[l] = (float*)(p + off_kd[l]);
        ix->knn_set_id[0][l] = ix->bb.knn_id[l]; ix->knn_set_d[0][l] = ix->bb.knn_d[l];
        for (int e = 0; e < 2; ++e) { ix->knn_set_id[e + 1][l] = (uint32_t*)(p + off_kis[e][l]); ix->knn_set_d[e + 1][l] = (float*)(p + off_kds[e][l]); }
    }
    ix->lt.n_edges = (uint32_t*)(p + o_ne); ix->lt.n_touched = ix->lt.n_edges + 1;
    ix->lt.edge_row = (uint32_t*)(p + o_er); ix->lt.edge_src = (uint32_t*)(p + o_es);
    ix->lt.touched_row = (uint32_t*)(p + o_tr); ix->lt.touched_slot = (uint32_t*)(p + o_ts); ix->lt.touched_level = (uint8_t*)(p + o_tl);
    ix->lt.inbox_fill = (uint32_t*)(p + o_if); ix->lt.inbox = (uint32_t*)(p + o_in);
    CU(ix->exact_scratch.ensure(exact_scratch_bytes_upto(B, KNN_BATCH, ix->dev.sm_count)));
    return VS_OK;
}

// K4 + K5 for slots [first, first + count): search on the frozen graph, intra-batch exact
// neighbours, heuristic selection, reverse-link merge.
// Levels and item ranges of the batch [first, first + count).
static void batch_shape(vs_index* ix, uint32_t first, uint32_t count, BuildBatch& b) {
    b.first_slot = first; b.count = count;
    b.new_top = top_for(ix, first + count);
    uint32_t off = 0;
    for (int l = 0; l < MAX_LEVELS; ++l) {
        b.item_off[l] = off;
        if (l <= b.new_top) {
            const uint32_t s = ix->stride[l];
            const uint32_t fr = (first + s - 1) / s, lr = (first + count - 1) / s;
            b.first_row[l] = fr; b.items[l] = lr >= fr ? lr - fr + 1 : 0;
        } else { b.first_row[l] = 0; b.items[l] = 0; }
        off += b.items[l];
    }
    b.item_off[MAX_LEVELS] = off; b.total_items = off;
}

// Exact intra-batch neighbours (K1/K2/K1f) of the batch on stream ks, into knn buffer set `buf`. They depend on the batch's
// own rows only, not on the graph, so they may run any time after the rows are in place.
vs_status batch_knn(vs_index* ix, uint32_t first, uint32_t count, int buf, cudaStream_t ks) {
    BuildBatch b = ix->bb;
    batch_shape(ix, first, count, b);
    for (int l = 0; l <= b.new_top; ++l) {
        if (b.items[l] < 2) continue;
        uint32_t* kid = ix->knn_set_id[buf][l]; float* kd = ix->knn_set_d[buf][l];
        const uint32_t s = ix->stride[l], fs = b.first_row[l] * s;
        ExactCorpus c{}; c.vecs = ix->vecs + (size_t)fs * ix->dpad; c.sqn = ix->sqn + fs; c.row_stride = (size_t)s * ix->dpad; c.row_step = s;
        c.first_slot = fs; c.n = b.items[l]; c.deleted = nullptr; c.slot_key = nullptr; c.n_slots = first + count;
        ExactQueries q{}; q.q = c.vecs; q.stride = c.row_stride; q.nq = b.items[l]; q.dims = ix->dims; q.dpad = ix->dpad; q.self_first = fs; q.self_step = s;
        const uint32_t k = std::min<uint32_t>(KNN_BATCH, b.items[l] - 1);
        ExactScratch sc{ix->exact_scratch.p, ix->exact_scratch.n};
        if (ix->build_tc && exact_tc_supported(c, q, k)) {   // big level-0 blocks: tensor-core scan + exact re-rank (same results)
            size_t want = exact_tc_scratch_bytes(q.nq, c.n, k + 1, ix->dpad, ix->dev.sm_count);
            if (want > ix->tc_scratch.n) {   // size it once for every batch shape up to max_batch: a cudaFree + cudaMalloc per growing batch cost up to 300 ms
                // (every tile count, not only powers of two: the plan's slice count is not monotonic in the batch size, and a
                // 30 000-row batch that outgrew the scratch sized for 32 768 rows stalled the pipeline for 19 ms)
                for (uint32_t bs = 4096; bs <= ix->max_batch; bs += 128) want = std::max(want, exact_tc_scratch_bytes(bs, bs, KNN_BATCH + 1, ix->dpad, ix->dev.sm_count));
                want = std::max(want, exact_tc_scratch_bytes(ix->max_batch, ix->max_batch, KNN_BATCH + 1, ix->dpad, ix->dev.sm_count));
                CU(cudaStreamSynchronize(ix->stream)); CU(cudaStreamSynchronize(ks));   // the scratch is about to move
            }
            CU(ix->tc_scratch.ensure(want));
            ExactScratch ts{ix->tc_scratch.p, ix->tc_scratch.n};
            CU(launch_exact_tc(c, q, ix->metric, k, KNN_BATCH, kid, nullptr, kd, ts, sc, ix->dev, ks, &ix->tc_build_stats));
            continue;
        }
        CU(launch_exact(c, q, ix->metric, k, KNN_BATCH, kid, nullptr, kd, sc, ix->dev, ks));
    }
    return VS_OK;
}

// K4a for slots [first, first + count): the batch's search on the frozen graph. `b` comes back filled for batch_link.
vs_status batch_search(vs_index* ix, uint32_t first, uint32_t count, int buf, BuildBatch& b) {
    b = ix->bb;   // scratch pointers
    batch_shape(ix, first, count, b);
    for (int l = 0; l < MAX_LEVELS; ++l) {
        if (l > b.new_top || b.items[l] < 2) { b.knn_id[l] = nullptr; b.knn_d[l] = nullptr; }
        else { b.knn_id[l] = ix->knn_set_id[buf][l]; b.knn_d[l] = ix->knn_set_d[buf][l]; }
    }
    auto mark = [&]() { if (ix->build_trace) { cudaEvent_t e; if (cudaEventCreate(&e) == cudaSuccess) { cudaEventRecord(e, ix->stream); ix->trace_ev.push_back(e); } } };
    mark();
    GraphView frozen = ix->view(first, b.new_top);
    CU(launch_build_search(frozen, b, ix->efA, ix->dev, ix->stats_on ? ix->counters : nullptr, ix->stream));
    mark();
    return VS_OK;
}

// K4b + K5: once the batch's exact neighbours are ready (`knn_ready`, null = already ordered on this stream), heuristic
// selection and reverse-link merge.
vs_status batch_link(vs_index* ix, const BuildBatch& b, cudaEvent_t knn_ready) {
    auto mark = [&]() { if (ix->build_trace) { cudaEvent_t e; if (cudaEventCreate(&e) == cudaSuccess) { cudaEventRecord(e, ix->stream); ix->trace_ev.push_back(e); } } };
    if (knn_ready) CU(cudaStreamWaitEvent(ix->stream, knn_ready, 0));
    mark();
    CU(cudaMemsetAsync(ix->lt.n_edges, 0, 8, ix->stream));
    GraphView linked = ix->view(b.first_slot + b.count, b.new_top);
    CU(launch_link(linked, b, ix->lt, ix->efpA, KNN_BATCH, ix->dev, ix->stats_on ? ix->counters : nullptr, ix->stream));
    mark();
    ix->st_build_batches++;
    return VS_OK;
}

// Lands every staged add: replace = tombstone the old slot + append a new one (usearch.rs:214-225).
vs_status flush_locked(vs_index* ix, const float* d_rows /* device rows or null => pend_vecs */, size_t n_dev, const uint64_t* dev_keys) {
    const bool from_dev = d_rows != nullptr;
    const size_t n = from_dev ? n_dev : ix->pend_keys.size();
    if (n == 0) return VS_OK;
    cudaSetDevice(ix->dev.device);
    const uint64_t* keys = from_dev ? dev_keys : ix->pend_keys.data();
    // capacity growth: the reference's threshold / increment rule (usearch.rs:62-66,201-212)
    while (ix->capacity - ix->n_slots < std::max(ix->reserve_increment / 3, n)) {
        vs_status s = do_reserve(ix, ix->capacity + std::max(ix->reserve_increment, n)); if (s != VS_OK) return s;
    }
    const uint32_t base = ix->n_slots;
    std::vector<uint32_t> dead;
    for (size_t i = 0; i < n; ++i) {
        const uint32_t slot = base + (uint32_t)i;
        auto it = ix->key_slot.find(keys[i]);
        if (it != ix->key_slot.end()) { dead.push_back(it->second); it->second = slot; ix->live--; }
        else ix->key_slot.emplace(keys[i], slot);
        ix->host_slot_key[slot] = keys[i];
    }
    if (!dead.empty()) {
        CU(ix->tomb.ensure(dead.size()));
        CU(cudaMemcpyAsync(ix->tomb.p, dead.data(), dead.size() * 4, cudaMemcpyHostToDevice, ix->stream));
        CU(launch_tombstone(ix->deleted, ix->tomb.p, (uint32_t)dead.size(), ix->stream));
    }
    CU(cudaMemcpyAsync(ix->slot_key + base, keys, n * 8, cudaMemcpyHostToDevice, ix->stream));
    if (!from_dev) {
        CU(ix->stage.ensure(n * ix->dims * 4));
        CU(cudaMemcpyAsync(ix->stage.p, ix->pend_vecs.data(), n * ix->dims * 4, cudaMemcpyHostToDevice, ix->stream));
        d_rows = (const float*)ix->stage.p;
    }
    CU(launch_row_prepare(d_rows, ix->vecs, ix->sqn, base, (uint32_t)n, ix->dims, ix->dpad, ix->stream));
    CU(cudaEventRecord(ix->ev0, ix->stream));
    // bootstrap block (all-pairs), then batches of at most 1/batch_div of the landed graph so
    // that almost every candidate a new node needs is already linked (recall vs batch size: DESIGN.md)
    std::vector<std::pair<uint32_t, uint32_t>> bl;
    for (size_t done = 0; done < n;) {
        const uint32_t cur = base + (uint32_t)done;
        uint32_t bsz = cur == 0 ? 1024u : std::max(256u, cur / ix->batch_div);
        bsz = (uint32_t)std::min<size_t>({(size_t)bsz, (size_t)ix->max_batch, n - done});
        bl.emplace_back(cur, bsz); done += bsz;
    }
    if (!ix->build_overlap) {            // one stream: neighbours, search, link, batch after batch
        for (auto& bt : bl) {
            BuildBatch b;
            vs_status s = batch_knn(ix, bt.first, bt.second, 0, ix->stream); if (s != VS_OK) return s;
            s = batch_search(ix, bt.first, bt.second, 0, b); if (s != VS_OK) return s;
            s = batch_link(ix, b, nullptr); if (s != VS_OK) return s;
        }
    } else {
        // The exact intra-batch neighbours (K1/K2/K1f) need the batch's rows only, so they run on the side stream - next to the
        // graph search of the same batch, and with build_pipeline = LA batches AHEAD: the neighbours of batch i+LA are enqueued while
        // batch i searches and links (the persistent search kernel fills every SM, the link kernels leave room). LA+1 buffer sets
        // rotate; a set is reused once the link step that read it has finished.
        const int LA = ix->build_pipeline, NS = LA + 1;             // look-ahead in batches, buffer sets in rotation
        bool linked_once[vs_index::KNN_SETS] = {};
        CU(cudaEventRecord(ix->fork_ev, ix->stream));               // rows are in place (launch_row_prepare above)
        CU(cudaStreamWaitEvent(ix->side, ix->fork_ev, 0));
        auto enqueue_knn = [&](size_t i) -> vs_status {
            const int buf = (int)(i % NS);
            if (linked_once[buf]) CU(cudaStreamWaitEvent(ix->side, ix->link_done[buf], 0));
            vs_status s = batch_knn(ix, bl[i].first, bl[i].second, buf, ix->side); if (s != VS_OK) return s;
            CU(cudaEventRecord(ix->knn_done[buf], ix->side));
            return VS_OK;
        };
        // Host order matters as much as stream order: a batch costs ~1 ms of host time to enqueue (the neighbour search alone is
        // a dozen launches, tensor maps and attribute calls per level) against ~5 ms on the GPU, so the graph search of batch i
        // goes to the GPU FIRST and the host prepares the side-stream work while it runs.
        for (size_t i = 0; i < (size_t)LA && i < bl.size(); ++i) { vs_status s = enqueue_knn(i); if (s != VS_OK) return s; }
        for (size_t i = 0; i < bl.size(); ++i) {
            const int buf = (int)(i % NS);
            BuildBatch b;
            vs_status s = batch_search(ix, bl[i].first, bl[i].second, buf, b); if (s != VS_OK) return s;
            if (i + LA < bl.size()) { s = enqueue_knn(i + LA); if (s != VS_OK) return s; }
            s = batch_link(ix, b, ix->knn_done[buf]); if (s != VS_OK) return s;
            CU(cudaEventRecord(ix->link_done[buf], ix->stream)); linked_once[buf] = true;
        }
    }
    CU(cudaEventRecord(ix->ev1, ix->stream));
    CU(cudaStreamSynchronize(ix->stream));
    float ms = 0; cudaEventElapsedTime(&ms, ix->ev0, ix->ev1); ix->st_build_seconds += ms * 1e-3;
    if (ix->build_trace && !ix->trace_ev.empty()) {   // per-phase device time and the gaps between batches
        double ph[3] = {0, 0, 0}, gaps = 0, worst = 0; size_t worst_i = 0;
        const size_t nb = ix->trace_ev.size() / 4;
        for (size_t i = 0; i < nb; ++i) {
            float t;
            for (int k = 0; k < 3; ++k) { cudaEventElapsedTime(&t, ix->trace_ev[4 * i + k], ix->trace_ev[4 * i + k + 1]); ph[k] += t; }
            cudaEventElapsedTime(&t, ix->trace_ev[4 * i], ix->trace_ev[4 * i + 3]);
            if (t > worst) { worst = t; worst_i = i; }
            if (i + 1 < nb) { cudaEventElapsedTime(&t, ix->trace_ev[4 * i + 3], ix->trace_ev[4 * i + 4]); gaps += t; }
        }
        float w3[3] = {0, 0, 0};
        for (int k = 0; k < 3 && nb; ++k) cudaEventElapsedTime(&w3[k], ix->trace_ev[4 * worst_i + k], ix->trace_ev[4 * worst_i + k + 1]);
        fprintf(stderr, "[vsann build trace] %zu batches: total %.1f ms = search %.1f + batch-knn %.1f + link %.1f + gaps %.1f; slowest batch #%zu %.2f ms (search %.2f, knn wait %.2f, link %.2f)\n",
                nb, ms, ph[0], ph[1], ph[2], gaps, worst_i, worst, w3[0], w3[1], w3[2]);
        for (cudaEvent_t e : ix->trace_ev) cudaEventDestroy(e);
        ix->trace_ev.clear();
    }
    ix->n_slots = base + (uint32_t)n; ix->live += n; ix->top_level = top_for(ix, ix->n_slots);
    ix->st_build_vectors += n;
    if (!from_dev) { ix->pend_keys.clear(); ix->pend_vecs.clear(); ix->pend_pos.clear(); }
    return VS_OK;
}

vs_status wait_user(vs_index* ix, cudaStream_t st) {
    if (ix->user_pending) { CU(cudaStreamWaitEvent(st, ix->user_ev, 0)); }
    return VS_OK;
}

// Search with device-resident queries/outputs on stream st (already ordered after the build).
vs_status search_device_locked(vs_index* ix, size_t nq, const float* d_q, size_t k, size_t ef, int exact,
                               uint64_t* d_keys, float* d_dist, cudaStream_t st) {
    unsigned long long* ctr = ix->stats_on ? ix->counters : nullptr;
    if (exact) {
        CU(ix->qpad.ensure(nq * ix->dpad));
        CU(ix->exact_scratch.ensure(exact_scratch_bytes((uint32_t)nq, (uint32_t)k, ix->dev.sm_count)));
        // pad the queries to dpad floats per row (aligned float4 loads in the tile kernel)
        CU(pad_queries(d_q, ix->qpad.p, (uint32_t)nq, ix->dims, ix->dpad, st));
        ExactCorpus c{}; c.vecs = ix->vecs; c.sqn = ix->sqn; c.row_stride = ix->dpad; c.row_step = 1; c.first_slot = 0; c.n = ix->n_slots;
        c.deleted = ix->deleted; c.slot_key = ix->slot_key; c.n_slots = ix->n_slots;
        ExactQueries q{}; q.q = ix->qpad.p; q.stride = ix->dpad; q.nq = (uint32_t)nq; q.dims = ix->dims; q.dpad = ix->dpad; q.self_first = 0; q.self_step = 0;
        ExactScratch sc{ix->exact_scratch.p, ix->exact_scratch.n};
        // K1s, the HBM-streaming scan: always for <= 4 queries (measured on B200, 5M x 1536 / 12.5M x 96 / 1.25M x 768: 4.4 / 0.9 / 0.7 ms
        // against 5.3 / 1.3 / 0.9 ms for the tensor path, whose fixed costs - sampling pass, re-rank, select - only pay off
        // from ~8 queries), up to 16 queries when the tensor path does not apply (k > 128, small corpora)
        const bool tc_ok = exact_tc_supported(c, q, (uint32_t)k);
        const bool stream = exact_stream_supported(c, q, (uint32_t)k, ix->dev) && (nq <= 4 || !tc_ok);
        if (stream) ix->st_stream_queries += nq;
        if (!stream && tc_ok) {
            // tcgen05 TF32 scan + exact FP32 re-rank (exact_tc.cu); same ids and distance bits as the FP32 kernel
            CU(ix->tc_scratch.ensure(exact_tc_scratch_bytes((uint32_t)nq, c.n, (uint32_t)k, ix->dpad, ix->dev.sm_count)));
            ExactScratch ts{ix->tc_scratch.p, ix->tc_scratch.n};
            CU(launch_exact_tc(c, q, ix->metric, (uint32_t)k, (uint32_t)k, nullptr, d_keys, d_dist, ts, sc, ix->dev, st, &ix->tc_stats));
            return VS_OK;
        }
        CU(launch_exact(c, q, ix->metric, (uint32_t)k, (uint32_t)k, nullptr, d_keys, d_dist, sc, ix->dev, st));
        return VS_OK;
    }
    const uint32_t efe = (uint32_t)std::max(ef ? ef : (size_t)ix->efS, k);
    GraphView g = ix->view(ix->n_slots, ix->top_level);
    cudaError_t e = launch_search(g, d_q, (uint32_t)nq, (uint32_t)k, efe, ix->dev, d_keys, d_dist, ctr, st);
    if (e == cudaErrorInvalidConfiguration) return fail(VS_ERR_UNSUPPORTED, "search: ef / dims too large for one warp's shared memory");
    CU(e);
    return VS_OK;
}

}  // namespace


// =============================================================== C ABI (include/vsann.h) ======
extern "C" {

vs_status vs_index_create(const vs_options* o, vs_index** out) {
    if (!o || !out) return fail(VS_ERR_INVALID, "create: null argument");
    *out = nullptr;
    if (o->dims == 0) return fail(VS_ERR_DIMS, "create: dimensions == 0");
    if (o->metric < 0 || o->metric > 2) return fail(VS_ERR_INVALID, "create: unknown metric");
    const size_t M = o->connectivity ? o->connectivity : 16;
    if (M < 2 || M > 16) return fail(VS_ERR_UNSUPPORTED, "create: connectivity must be in [2, 16] (level-0 row = 2M <= 32 links, one per lane)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(VS_ERR_CUDA, "create: no CUDA device (libvsann has no CPU fallback)"); }
    if (o->device < 0 || o->device >= ndev) return fail(VS_ERR_INVALID, "create: bad device ordinal");
    cudaDeviceProp prop{};
    CU(cudaGetDeviceProperties(&prop, o->device));
    if (prop.major < 10) return fail(VS_ERR_CUDA, "create: device is not Blackwell (sm_100a kernels only)");
    CU(cudaSetDevice(o->device));
    auto* ix = new vs_index();
    ix->dims = (uint32_t)o->dims; ix->dpad = (ix->dims + 3u) & ~3u; ix->metric = o->metric;
    ix->M = (uint32_t)M; ix->W0 = 2 * ix->M;
    ix->efA = (uint32_t)(o->expansion_add ? o->expansion_add : 128);
    ix->efS = (uint32_t)(o->expansion_search ? o->expansion_search : 64);
    ix->max_batch = (uint32_t)(o->max_batch ? o->max_batch : 32768);
    if (const char* e = getenv("VSANN_MAX_BATCH")) { const long v = atol(e); if (v > 0) ix->max_batch = (uint32_t)v; }
    if (ix->max_batch < 32) ix->max_batch = 32;
    if (const char* e = getenv("VSANN_BATCH_DIV")) { const long v = atol(e); if (v > 0) ix->batch_div = (uint32_t)v; }
    ix->reserve_increment = o->capacity_hint ? o->capacity_hint : 1000000;   // usearch.rs:62
    ix->dev.device = o->device; ix->dev.sm_count = prop.multiProcessorCount; ix->dev.smem_per_block = prop.sharedMemPerBlockOptin;
    if (const char* e = getenv("VSANN_HASH_FACTOR")) { const long v = atol(e); if (v > 0) ix->dev.hash_factor = (uint32_t)v; }
    if (const char* e = getenv("VSANN_HASH_MIN")) { const long v = atol(e); if (v >= 64) ix->dev.hash_min = (uint32_t)v; }
    if (const char* e = getenv("VSANN_SMEM_LIST")) ix->dev.force_smem_list = atoi(e) != 0;
    if (const char* e = getenv("VSANN_STAGE_BYTES")) ix->dev.stage_bytes = (uint32_t)atol(e);
    if (const char* e = getenv("VSANN_GATHER4")) ix->dev.gather4 = atoi(e) != 0;
    if (const char* e = getenv("VSANN_SELECT_SMEM")) ix->dev.select_smem = atoi(e) != 0;
    if (const char* e = getenv("VSANN_ZERO_COPY")) ix->zero_copy = atoi(e) != 0;   // 0: always stage host queries with a copy
    if (const char* e = getenv("VSANN_COALESCE")) ix->coalesce = atoi(e) != 0;     // 0: every vs_search call is its own launch
    if (const char* e = getenv("VSANN_BUILD_PIPELINE")) ix->build_pipeline = std::max(0, std::min(vs_index::KNN_SETS - 1, atoi(e)));   // batches of look-ahead; 0: a batch's neighbours next to its own graph search only
    if (const char* e = getenv("VSANN_BUILD_OVERLAP")) ix->build_overlap = atoi(e) != 0;   // 0: batch neighbours after the graph search, same stream
    if (const char* e = getenv("VSANN_BUILD_TRACE")) ix->build_trace = atoi(e) != 0;   // per-phase build timing on stderr
    if (const char* e = getenv("VSANN_BUILD_TC")) ix->build_tc = atoi(e) != 0;   // 0: FP32 tile kernel for the batch neighbours (A/B, profiling)
    vs_status s = VS_OK;
    auto body = [&]() -> vs_status {
        CU(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&ix->side, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&ix->fork_ev, cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&ix->join_ev, cudaEventDisableTiming));
        for (int e = 0; e < vs_index::KNN_SETS; ++e) { CU(cudaEventCreateWithFlags(&ix->knn_done[e], cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&ix->link_done[e], cudaEventDisableTiming)); }
        CU(cudaEventCreate(&ix->ev0)); CU(cudaEventCreate(&ix->ev1)); CU(cudaEventCreateWithFlags(&ix->user_ev, cudaEventDisableTiming));
        CU(cudaMalloc(&ix->counters, 8 * sizeof(unsigned long long)));
        CU(cudaMemsetAsync(ix->counters, 0, 8 * sizeof(unsigned long long), ix->stream));
        CU(cudaMalloc(&ix->dev.work_ring, WORK_RING_SLOTS * sizeof(uint32_t)));
        vs_status r = do_reserve(ix, ix->reserve_increment); if (r != VS_OK) return r;
        r = setup_build_scratch(ix); if (r != VS_OK) return r;
        CU(cudaStreamSynchronize(ix->stream));
        return VS_OK;
    };
    s = body();
    if (s != VS_OK) { std::string keep = g_err; vs_index_destroy(ix); g_err = keep; return s; }
    *out = ix; return VS_OK;
}

void vs_index_destroy(vs_index* ix) {
    if (!ix) return;
    cudaSetDevice(ix->dev.device);
    if (ix->stream) cudaStreamSynchronize(ix->stream);
    cudaFree(ix->vecs); cudaFree(ix->sqn); cudaFree(ix->links0); cudaFree(ix->slot_key); cudaFree(ix->deleted);
    for (int l = 1; l < MAX_LEVELS; ++l) cudaFree(ix->linksU[l]);
    cudaFree(ix->lt.inbox_cnt); cudaFree(ix->lt.inbox_idx); cudaFree(ix->counters); cudaFree(ix->dev.work_ring);
    ix->build_scratch.release(); ix->exact_scratch.release(); ix->tc_scratch.release(); ix->stage.release(); ix->tomb.release();
    ix->qpad.release(); ix->qdense.release(); ix->out_d.release(); ix->out_k.release();
    if (ix->ev0) cudaEventDestroy(ix->ev0); if (ix->ev1) cudaEventDestroy(ix->ev1); if (ix->user_ev) cudaEventDestroy(ix->user_ev);
    if (ix->fork_ev) cudaEventDestroy(ix->fork_ev); if (ix->join_ev) cudaEventDestroy(ix->join_ev);
    for (int e = 0; e < vs_index::KNN_SETS; ++e) { if (ix->knn_done[e]) cudaEventDestroy(ix->knn_done[e]); if (ix->link_done[e]) cudaEventDestroy(ix->link_done[e]); }
    if (ix->side) { cudaStreamSynchronize(ix->side); cudaStreamDestroy(ix->side); }
    if (ix->stream) cudaStreamDestroy(ix->stream);
    if (ix->found_pin) cudaFreeHost(ix->found_pin); ix->found_d.release();
    if (ix->co_pin_q) cudaFreeHost(ix->co_pin_q); if (ix->co_pin_k) cudaFreeHost(ix->co_pin_k); if (ix->co_pin_d) cudaFreeHost(ix->co_pin_d);
    cudaGetLastError();
    delete ix;
}

vs_status vs_add_batch(vs_index* ix, size_t n, const uint64_t* keys, const float* vecs) {
    if (!ix || (n && (!keys || !vecs))) return fail(VS_ERR_INVALID, "add: null argument");
    std::lock_guard<std::mutex> g(ix->mu);
    for (size_t i = 0; i < n; ++i) {
        auto it = ix->pend_pos.find(keys[i]);
        if (it != ix->pend_pos.end()) {   // replaced before it landed: overwrite in place
            std::memcpy(ix->pend_vecs.data() + it->second * ix->dims, vecs + i * ix->dims, ix->dims * sizeof(float));
            continue;
        }
        ix->pend_pos.emplace(keys[i], ix->pend_keys.size());
        ix->pend_keys.push_back(keys[i]);
        ix->pend_vecs.insert(ix->pend_vecs.end(), vecs + i * ix->dims, vecs + (i + 1) * ix->dims);
    }
    if (ix->pend_keys.size() >= ix->max_batch) return flush_locked(ix, nullptr, 0, nullptr);
    return VS_OK;
}
vs_status vs_add(vs_index* ix, uint64_t key, const float* vec) { return vs_add_batch(ix, 1, &key, vec); }

vs_status vs_add_batch_device(vs_index* ix, size_t n, const uint64_t* keys, const float* d_vecs) {
    if (!ix || (n && (!keys || !d_vecs))) return fail(VS_ERR_INVALID, "add: null argument");
    std::lock_guard<std::mutex> g(ix->mu);
    vs_status s = flush_locked(ix, nullptr, 0, nullptr); if (s != VS_OK) return s;
    {   // keys must be distinct inside one device batch
        std::unordered_map<uint64_t, int> seen; seen.reserve(n * 2);
        for (size_t i = 0; i < n; ++i) if (!seen.emplace(keys[i], 1).second) return fail(VS_ERR_INVALID, "add_batch_device: duplicate key inside the batch");
    }
    CU(cudaStreamSynchronize(nullptr));   // rows were produced on the caller's (legacy default) stream or earlier
    return flush_locked(ix, d_vecs, n, keys);
}

vs_status vs_remove(vs_index* ix, uint64_t key) {
    if (!ix) return fail(VS_ERR_INVALID, "remove: null index");
    std::lock_guard<std::mutex> g(ix->mu);
    vs_status s = flush_locked(ix, nullptr, 0, nullptr); if (s != VS_OK) return s;
    auto it = ix->key_slot.find(key);
    if (it == ix->key_slot.end()) return fail(VS_NOT_FOUND, "remove: key not found");
    cudaSetDevice(ix->dev.device);
    const uint32_t slot = it->second;
    CU(ix->tomb.ensure(1));
    CU(cudaMemcpyAsync(ix->tomb.p, &slot, 4, cudaMemcpyHostToDevice, ix->stream));
    CU(launch_tombstone(ix->deleted, ix->tomb.p, 1, ix->stream));
    CU(cudaStreamSynchronize(ix->stream));
    ix->key_slot.erase(it); ix->live--;
    return VS_OK;
}

vs_status vs_search_batch_device(vs_index* ix, size_t nq, const float* d_q, size_t k, size_t ef, int exact,
                                 uint64_t* d_keys, float* d_dist, void* stream) {
    if (!ix || (nq && (!d_q || !d_keys || !d_dist))) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    if (nq == 0) return VS_OK;
    std::lock_guard<std::mutex> g(ix->mu);
    vs_status s = flush_locked(ix, nullptr, 0, nullptr); if (s != VS_OK) return s;
    cudaSetDevice(ix->dev.device);
    cudaStream_t st = stream ? (cudaStream_t)stream : ix->stream;
    s = wait_user(ix, st); if (s != VS_OK) return s;
    s = search_device_locked(ix, nq, d_q, k, ef, exact, d_keys, d_dist, st); if (s != VS_OK) return s;
    if (stream) { CU(cudaEventRecord(ix->user_ev, st)); ix->user_pending = true; }
    else { CU(cudaStreamSynchronize(st)); }
    return VS_OK;
}

vs_status vs_search_batch(vs_index* ix, size_t nq, const float* qs, size_t k, size_t ef, int exact,
                          uint64_t* keys_out, float* dist_out, size_t* found) {
    if (!ix || (nq && (!qs || !keys_out || !dist_out))) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    if (nq == 0) return VS_OK;
    std::lock_guard<std::mutex> g(ix->mu);
    vs_status s = flush_locked(ix, nullptr, 0, nullptr); if (s != VS_OK) return s;
    cudaSetDevice(ix->dev.device);
    CU(ix->out_k.ensure(nq * k)); CU(ix->out_d.ensure(nq * k));
    s = wait_user(ix, ix->stream); if (s != VS_OK) return s;
    // Pinned (page-locked, UVA-mapped) query buffers are read by the traversal kernel in place: every warp
    // fetches its 4 d bytes over PCIe while the others compute, so the H2D copy leaves the critical path.
    // Pageable buffers, and the exact path (which streams the queries many times through TMA), are staged.
    const float* d_q = nullptr;
    if (!exact && ix->zero_copy) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, qs) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) d_q = (const float*)at.devicePointer;
        else cudaGetLastError();
    }
    if (!d_q) {
        CU(ix->qdense.ensure(nq * ix->dims));
        CU(cudaMemcpyAsync(ix->qdense.p, qs, nq * ix->dims * 4, cudaMemcpyHostToDevice, ix->stream));
        d_q = ix->qdense.p;
    }
    ix->st_search_launches++; ix->st_search_calls += nq;
    s = search_device_locked(ix, nq, d_q, k, ef, exact, ix->out_k.p, ix->out_d.p, ix->stream); if (s != VS_OK) return s;
    // (results: writing them straight into pinned host memory from the kernel was measured and is no faster than
    // this 12 k B per query copy, so the explicit copy stays)
    CU(cudaMemcpyAsync(keys_out, ix->out_k.p, nq * k * 8, cudaMemcpyDeviceToHost, ix->stream));
    CU(cudaMemcpyAsync(dist_out, ix->out_d.p, nq * k * 4, cudaMemcpyDeviceToHost, ix->stream));
    if (found) {   // counted on the device, 4 bytes per query come back
        CU(ix->found_d.ensure(nq));
        if (nq > ix->found_cap) {
            if (ix->found_pin) cudaFreeHost(ix->found_pin);
            ix->found_pin = nullptr; ix->found_cap = 0;
            const size_t want = std::max<size_t>(nq, 16384);
            CU(cudaHostAlloc((void**)&ix->found_pin, want * 4, cudaHostAllocDefault));
            ix->found_cap = want;
        }
        CU(launch_count_found(ix->out_k.p, (uint32_t)nq, (uint32_t)k, ix->found_d.p, ix->stream));
        CU(cudaMemcpyAsync(ix->found_pin, ix->found_d.p, nq * 4, cudaMemcpyDeviceToHost, ix->stream));
    }
    CU(cudaStreamSynchronize(ix->stream));
    if (found) for (size_t i = 0; i < nq; ++i) found[i] = ix->found_pin[i];
    return VS_OK;
}

// One launch for every request of the round (requests with a different k go in separate launches). The queries are packed
// into a pinned staging buffer the traversal kernel reads in place; results come back through pinned buffers.
static void co_run(vs_index* ix, std::vector<vs_index::SearchReq*>& batch) {
    size_t done = 0;
    while (done < batch.size()) {
        const size_t k = batch[done]->k;
        std::vector<vs_index::SearchReq*> grp;
        for (size_t i = done; i < batch.size(); ++i) if (batch[i]->k == k && batch[i]->st == (vs_status)-1) grp.push_back(batch[i]);
        const size_t n = grp.size(), d = ix->dims;
        vs_status s = VS_OK;
        if (n * d > ix->co_cap_q) {
            if (ix->co_pin_q) cudaFreeHost(ix->co_pin_q);
            ix->co_pin_q = nullptr; ix->co_cap_q = 0;
            const size_t want = std::max<size_t>(n * d, 256 * d);
            if (cudaHostAlloc((void**)&ix->co_pin_q, want * 4, cudaHostAllocDefault) == cudaSuccess) ix->co_cap_q = want;
            else { cudaGetLastError(); s = fail(VS_ERR_NOMEM, "search: pinned staging allocation failed"); }
        }
        if (s == VS_OK && n * k > ix->co_cap_o) {
            if (ix->co_pin_k) cudaFreeHost(ix->co_pin_k); if (ix->co_pin_d) cudaFreeHost(ix->co_pin_d);
            ix->co_pin_k = nullptr; ix->co_pin_d = nullptr; ix->co_cap_o = 0;
            const size_t want = std::max<size_t>(n * k, 256 * k);
            if (cudaHostAlloc((void**)&ix->co_pin_k, want * 8, cudaHostAllocDefault) == cudaSuccess &&
                cudaHostAlloc((void**)&ix->co_pin_d, want * 4, cudaHostAllocDefault) == cudaSuccess) ix->co_cap_o = want;
            else { cudaGetLastError(); s = fail(VS_ERR_NOMEM, "search: pinned staging allocation failed"); }
        }
        if (s == VS_OK) {
            for (size_t i = 0; i < n; ++i) std::memcpy(ix->co_pin_q + i * d, grp[i]->q, d * sizeof(float));
            s = vs_search_batch(ix, n, ix->co_pin_q, k, 0, 0, ix->co_pin_k, ix->co_pin_d, nullptr);
        }
        for (size_t i = 0; i < n; ++i) {
            vs_index::SearchReq* r = grp[i];
            r->st = s;
            if (s != VS_OK) { r->err = g_err; continue; }
            std::memcpy(r->keys, ix->co_pin_k + i * k, k * 8); std::memcpy(r->dist, ix->co_pin_d + i * k, k * 4);
            if (r->found) { size_t f = 0; while (f < k && r->keys[f] != UINT64_MAX) ++f; *r->found = f; }
        }
        while (done < batch.size() && batch[done]->st != (vs_status)-1) ++done;
    }
}

// usearch.rs:276 is called from up to 2 x T tasks at once (usearch.rs:118-131), each with ONE query. A GPU launch per query
// would serialise them at ~0.3 ms each, so concurrent callers are coalesced: the first caller to arrive becomes the leader
// and launches at once (no added latency on an idle index); whoever arrives while that launch is in flight queues up, and
// the next leader takes the whole queue in one launch. Batch size adapts to the load by itself, no timer involved.
vs_status vs_search(vs_index* ix, const float* q, size_t k, uint64_t* keys_out, float* dist_out, size_t* found) {
    if (!ix || !q || !keys_out || !dist_out) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    if (!ix->coalesce) return vs_search_batch(ix, 1, q, k, 0, 0, keys_out, dist_out, found);
    vs_index::SearchReq r{q, k, keys_out, dist_out, found, (vs_status)-1, std::string(), false};
    std::unique_lock<std::mutex> l(ix->co_mu);
    ix->co_q.push_back(&r);
    while (!r.done) {
        if (!ix->co_leader) {
            ix->co_leader = true;
            std::vector<vs_index::SearchReq*> batch; batch.swap(ix->co_q);
            l.unlock();
            co_run(ix, batch);
            l.lock();
            for (vs_index::SearchReq* b : batch) b->done = true;
            ix->co_leader = false;
            ix->co_cv.notify_all();
        } else {
            ix->co_cv.wait(l);
        }
    }
    l.unlock();
    if (r.st != VS_OK) g_err = r.err;
    return r.st;
}

size_t vs_count(vs_index* ix) {
    if (!ix) return 0;
    std::lock_guard<std::mutex> g(ix->mu);
    flush_locked(ix, nullptr, 0, nullptr);
    return ix->live;
}
size_t vs_capacity(vs_index* ix) { if (!ix) return 0; std::lock_guard<std::mutex> g(ix->mu); return ix->capacity; }
vs_status vs_reserve(vs_index* ix, size_t cap) {
    if (!ix) return fail(VS_ERR_INVALID, "reserve: null index");
    std::lock_guard<std::mutex> g(ix->mu); return do_reserve(ix, cap);
}
vs_status vs_flush(vs_index* ix) {
    if (!ix) return fail(VS_ERR_INVALID, "flush: null index");
    std::lock_guard<std::mutex> g(ix->mu); return flush_locked(ix, nullptr, 0, nullptr);
}

vs_status vs_get_stats(vs_index* ix, vs_stats* o) {
    if (!ix || !o) return fail(VS_ERR_INVALID, "stats: null argument");
    std::lock_guard<std::mutex> g(ix->mu);
    std::memset(o, 0, sizeof(*o));
    unsigned long long c[8];
    cudaSetDevice(ix->dev.device);
    CU(cudaStreamSynchronize(ix->stream));
    CU(cudaMemcpy(c, ix->counters, sizeof(c), cudaMemcpyDeviceToHost));
    o->count = ix->live; o->slots = ix->n_slots; o->capacity = ix->capacity; o->pending = ix->pend_keys.size();
    o->search_queries = c[0]; o->search_dist_evals = c[1]; o->search_expansions = c[2];
    o->build_vectors = ix->st_build_vectors; o->build_dist_evals = c[3]; o->build_batches = ix->st_build_batches;
    o->build_seconds = ix->st_build_seconds; o->device_bytes = ix->device_bytes + ix->build_scratch.n + ix->exact_scratch.n;
    o->top_level = (uint32_t)ix->top_level; o->dims = ix->dims;
    o->exact_tc_queries = ix->tc_stats.queries; o->exact_tc_fallback = ix->tc_stats.fp32_fallback; o->exact_tc_second_scan = ix->tc_stats.second_scan;
    o->exact_tc_scans = ix->tc_stats.scans; o->exact_tc_scan_seconds = ix->tc_stats.scan_ms * 1e-3;
    o->exact_stream_queries = ix->st_stream_queries; o->search_host_queries = ix->st_search_calls; o->search_host_launches = ix->st_search_launches;
    return VS_OK;
}
vs_status vs_set_stats(vs_index* ix, int on) { if (!ix) return VS_ERR_INVALID; std::lock_guard<std::mutex> g(ix->mu); ix->stats_on = on != 0; ix->tc_stats.time_scan = on != 0; return VS_OK; }
vs_status vs_reset_stats(vs_index* ix) {
    if (!ix) return VS_ERR_INVALID;
    std::lock_guard<std::mutex> g(ix->mu);
    cudaSetDevice(ix->dev.device);
    CU(cudaMemsetAsync(ix->counters, 0, 8 * sizeof(unsigned long long), ix->stream));
    CU(cudaStreamSynchronize(ix->stream));
    ix->st_build_vectors = ix->st_build_batches = 0; ix->st_build_seconds = 0; ix->st_stream_queries = 0; ix->st_search_calls = ix->st_search_launches = 0; { const bool t = ix->tc_stats.time_scan; ix->tc_stats = TcStats(); ix->tc_stats.time_scan = t; }
    return VS_OK;
}

vs_status vs_merge_topk_device(int device, size_t parts, size_t nq, size_t k, const uint64_t* d_keys_in, const float* d_dist_in,
                               uint64_t* d_keys_out, float* d_dist_out, void* stream) {
    if (!d_keys_in || !d_dist_in || !d_keys_out || !d_dist_out) return fail(VS_ERR_INVALID, "merge: null argument");
    CU(cudaSetDevice(device));
    CU(launch_merge_topk((uint32_t)parts, (uint32_t)nq, (uint32_t)k, d_keys_in, d_dist_in, d_keys_out, d_dist_out, (cudaStream_t)stream));
    if (!stream) CU(cudaStreamSynchronize(nullptr));
    return VS_OK;
}

vs_status vs_merge_topk_peers_device(int device, size_t parts, size_t nq, size_t k, const uint64_t* const* d_part_keys, const float* const* d_part_dist,
                                     uint64_t* d_keys_out, float* d_dist_out, void* stream) {
    if (!d_part_keys || !d_part_dist || !d_keys_out || !d_dist_out) return fail(VS_ERR_INVALID, "merge: null argument");
    CU(cudaSetDevice(device));
    CU(launch_merge_topk_ptrs((uint32_t)parts, (uint32_t)nq, (uint32_t)k, d_part_keys, d_part_dist, d_keys_out, d_dist_out, (cudaStream_t)stream));
    if (!stream) CU(cudaStreamSynchronize(nullptr));
    return VS_OK;
}

vs_status vs_export_level(vs_index* ix, uint32_t level, uint32_t* links_out, size_t* n_rows, size_t* row_width, uint64_t* slot_keys_out, size_t* slots) {
    if (!ix || !n_rows || !row_width) return fail(VS_ERR_INVALID, "export: null argument");
    std::lock_guard<std::mutex> g(ix->mu);
    vs_status s = flush_locked(ix, nullptr, 0, nullptr); if (s != VS_OK) return s;
    if ((int)level >= ix->n_levels) return fail(VS_ERR_INVALID, "export: no such level");
    const size_t st = ix->stride[level], nr = ix->n_slots ? (ix->n_slots - 1) / st + 1 : 0, w = level == 0 ? ix->W0 : ix->M;
    *n_rows = nr; *row_width = w; if (slots) *slots = ix->n_slots;
    cudaSetDevice(ix->dev.device);
    if (links_out && nr) CU(cudaMemcpy(links_out, level == 0 ? ix->links0 : ix->linksU[level], nr * w * 4, cudaMemcpyDeviceToHost));
    if (slot_keys_out && ix->n_slots) CU(cudaMemcpy(slot_keys_out, ix->slot_key, (size_t)ix->n_slots * 8, cudaMemcpyDeviceToHost));
    return VS_OK;
}

// ---- persistence (vsann.h; the reference rebuilds from a table scan instead, db_index.rs:108-117,217-237) ----
namespace {
struct FileHeader {          // 128 bytes, little endian
    char magic[8];           // "VSANNIDX"
    uint32_t version, dims, dpad, metric, M, efA, efS, n_levels_used;
    uint64_t n_slots, live;
    uint32_t max_batch, reserved32;
    uint64_t reserved[8];
};
static_assert(sizeof(FileHeader) == 128, "header layout");
constexpr uint32_t FILE_VERSION = 1;

struct Summer {              // order-dependent 64-bit checksum over 8-byte words (tail bytes zero-padded)
    uint64_t h = 0x9E3779B97F4A7C15ull;
    void add(const void* p, size_t n) {
        const unsigned char* b = (const unsigned char*)p;
        size_t i = 0;
        for (; i + 8 <= n; i += 8) { uint64_t w; std::memcpy(&w, b + i, 8); h = (h ^ w) * 0x100000001B3ull; h ^= h >> 29; }
        if (i < n) { uint64_t w = 0; std::memcpy(&w, b + i, n - i); h = (h ^ w) * 0x100000001B3ull; h ^= h >> 29; }
    }
};

// device array <-> file through a bounded host buffer
vs_status dump(FILE* f, Summer& sum, const void* dptr, size_t bytes, std::vector<unsigned char>& buf) {
    for (size_t off = 0; off < bytes; off += buf.size()) {
        const size_t n = std::min(buf.size(), bytes - off);
        CU(cudaMemcpy(buf.data(), (const unsigned char*)dptr + off, n, cudaMemcpyDeviceToHost));
        sum.add(buf.data(), n);
        if (fwrite(buf.data(), 1, n, f) != n) return fail(VS_ERR_INVALID, "save: short write");
    }
    return VS_OK;
}
vs_status restore(FILE* f, Summer& sum, void* dptr, size_t bytes, std::vector<unsigned char>& buf) {
    for (size_t off = 0; off < bytes; off += buf.size()) {
        const size_t n = std::min(buf.size(), bytes - off);
        if (fread(buf.data(), 1, n, f) != n) return fail(VS_ERR_INVALID, "load: file is truncated");
        sum.add(buf.data(), n);
        CU(cudaMemcpy((unsigned char*)dptr + off, buf.data(), n, cudaMemcpyHostToDevice));
    }
    return VS_OK;
}
struct FileCloser { FILE* f; ~FileCloser() { if (f) fclose(f); } };
}  // namespace

vs_status vs_index_save(vs_index* ix, const char* path) {
    if (!ix || !path) return fail(VS_ERR_INVALID, "save: null argument");
    std::lock_guard<std::mutex> g(ix->mu);
    vs_status s = flush_locked(ix, nullptr, 0, nullptr); if (s != VS_OK) return s;
    cudaSetDevice(ix->dev.device);
    CU(cudaStreamSynchronize(ix->stream));
    FILE* f = fopen(path, "wb");
    if (!f) return fail(VS_ERR_INVALID, std::string("save: cannot open ") + path);
    FileCloser fc{f};
    FileHeader h{};
    std::memcpy(h.magic, "VSANNIDX", 8);
    h.version = FILE_VERSION; h.dims = ix->dims; h.dpad = ix->dpad; h.metric = (uint32_t)ix->metric; h.M = ix->M; h.efA = ix->efA; h.efS = ix->efS;
    h.n_levels_used = (uint32_t)top_for(ix, ix->n_slots) + 1; h.n_slots = ix->n_slots; h.live = ix->live; h.max_batch = ix->max_batch;
    Summer sum; sum.add(&h, sizeof h);
    if (fwrite(&h, sizeof h, 1, f) != 1) return fail(VS_ERR_INVALID, "save: short write");
    std::vector<unsigned char> buf(64u << 20);
    const size_t n = ix->n_slots;
    if ((s = dump(f, sum, ix->vecs, n * ix->dpad * 4, buf)) != VS_OK) return s;
    if ((s = dump(f, sum, ix->sqn, n * 4, buf)) != VS_OK) return s;
    if ((s = dump(f, sum, ix->links0, n * ix->W0 * 4, buf)) != VS_OK) return s;
    for (uint32_t l = 1; l < h.n_levels_used; ++l) {
        const size_t nr = (n + ix->stride[l] - 1) / ix->stride[l];
        if ((s = dump(f, sum, ix->linksU[l], nr * ix->M * 4, buf)) != VS_OK) return s;
    }
    if ((s = dump(f, sum, ix->slot_key, n * 8, buf)) != VS_OK) return s;
    if ((s = dump(f, sum, ix->deleted, ((n + 31) / 32) * 4, buf)) != VS_OK) return s;
    if (fwrite(&sum.h, 8, 1, f) != 1 || fflush(f) != 0) return fail(VS_ERR_INVALID, "save: short write");
    return VS_OK;
}

vs_status vs_index_load(const char* path, const vs_options* ov, vs_index** out) {
    if (!path || !out) return fail(VS_ERR_INVALID, "load: null argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(VS_ERR_INVALID, std::string("load: cannot open ") + path);
    FileCloser fc{f};
    FileHeader h{};
    if (fread(&h, sizeof h, 1, f) != 1 || std::memcmp(h.magic, "VSANNIDX", 8) != 0) return fail(VS_ERR_INVALID, "load: not a vsann index file");
    if (h.version > FILE_VERSION) return fail(VS_ERR_UNSUPPORTED, "load: file format " + std::to_string(h.version) + " is newer than this library");
    if (h.version == 0 || h.dims == 0 || h.dpad != ((h.dims + 3u) & ~3u) || h.metric > 2 || h.M < 2 || h.M > 16 || h.n_slots >= 0x7FFFFFF0ull || h.live > h.n_slots)
        return fail(VS_ERR_INVALID, "load: corrupt header");
    vs_options o{};
    o.dims = h.dims; o.metric = (int)h.metric; o.connectivity = h.M; o.expansion_add = h.efA; o.expansion_search = h.efS;
    o.device = ov ? ov->device : 0; o.max_batch = ov && ov->max_batch ? ov->max_batch : h.max_batch;
    const size_t hint = ov && ov->capacity_hint ? ov->capacity_hint : 1000000;
    o.capacity_hint = hint;
    vs_index* ix = nullptr;
    vs_status s = vs_index_create(&o, &ix); if (s != VS_OK) return s;
    auto body = [&]() -> vs_status {
        std::lock_guard<std::mutex> g(ix->mu);
        const size_t n = (size_t)h.n_slots;
        vs_status r = do_reserve(ix, n + std::max<size_t>(hint / 3, 1)); if (r != VS_OK) return r;
        if ((uint32_t)top_for(ix, (uint32_t)n) + 1 != h.n_levels_used) return fail(VS_ERR_INVALID, "load: level count does not match the slot count");
        Summer sum; sum.add(&h, sizeof h);
        std::vector<unsigned char> buf(64u << 20);
        if ((r = restore(f, sum, ix->vecs, n * ix->dpad * 4, buf)) != VS_OK) return r;
        if ((r = restore(f, sum, ix->sqn, n * 4, buf)) != VS_OK) return r;
        if ((r = restore(f, sum, ix->links0, n * ix->W0 * 4, buf)) != VS_OK) return r;
        for (uint32_t l = 1; l < h.n_levels_used; ++l) {
            const size_t nr = (n + ix->stride[l] - 1) / ix->stride[l];
            if ((r = restore(f, sum, ix->linksU[l], nr * ix->M * 4, buf)) != VS_OK) return r;
        }
        // slot keys and tombstones are also needed on the host: key -> newest live slot
        std::vector<uint64_t> keys(n); std::vector<uint32_t> del((n + 31) / 32);
        if (n && fread(keys.data(), 8, n, f) != n) return fail(VS_ERR_INVALID, "load: file is truncated");
        sum.add(keys.data(), n * 8);
        if (n && fread(del.data(), 4, del.size(), f) != del.size()) return fail(VS_ERR_INVALID, "load: file is truncated");
        sum.add(del.data(), del.size() * 4);
        uint64_t want = 0;
        if (fread(&want, 8, 1, f) != 1) return fail(VS_ERR_INVALID, "load: file is truncated");
        if (want != sum.h) return fail(VS_ERR_INVALID, "load: checksum mismatch");
        if (n) { CU(cudaMemcpy(ix->slot_key, keys.data(), n * 8, cudaMemcpyHostToDevice)); CU(cudaMemcpy(ix->deleted, del.data(), del.size() * 4, cudaMemcpyHostToDevice)); }
        size_t live = 0;
        ix->key_slot.reserve(h.live * 2);
        for (size_t sl = 0; sl < n; ++sl) {
            ix->host_slot_key[sl] = keys[sl];
            if (del[sl >> 5] >> (sl & 31) & 1u) continue;
            if (!ix->key_slot.emplace(keys[sl], (uint32_t)sl).second) return fail(VS_ERR_INVALID, "load: two live slots share a key");
            ++live;
        }
        if (live != h.live) return fail(VS_ERR_INVALID, "load: live count does not match the tombstones");
        ix->n_slots = (uint32_t)n; ix->live = live; ix->top_level = top_for(ix, ix->n_slots);
        return VS_OK;
    };
    s = body();
    if (s != VS_OK) { std::string keep = g_err; vs_index_destroy(ix); g_err = keep; return s; }
    *out = ix; return VS_OK;
}

const char* vs_last_error(void) { return g_err.c_str(); }
const char* vs_version(void) { return "vsann-b200 0.1 sm_100a"; }

}  // extern "C"
