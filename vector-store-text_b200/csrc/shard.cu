// shard.cu — one GPU's part of an index (see shard.cuh). There is no CPU path: every operation that touches vectors runs on the GPU.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "shard.cuh"

namespace vsann {

thread_local std::string g_err;
vs_status fail(vs_status s, const std::string& m) { g_err = m; return s; }

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

__global__ void k_pad(const float* __restrict__ src, float* __restrict__ dst, uint32_t n, uint32_t dims, uint32_t dpad) {
    const size_t total = (size_t)n * dpad;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(i / dpad), c = (uint32_t)(i % dpad);
        dst[i] = c < dims ? src[(size_t)r * dims + c] : 0.f;
    }
}
static cudaError_t pad_queries(const float* src, float* dst, uint32_t n, uint32_t dims, uint32_t dpad, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const size_t total = (size_t)n * dpad;
    uint32_t blocks = (uint32_t)std::min<size_t>((total + 255) / 256, 148 * 8);
    k_pad<<<blocks, 256, 0, st>>>(src, dst, n, dims, dpad);
    return cudaGetLastError();
}

// compaction: dst row i = src row map[i], rows of `words` 4-byte words
__global__ void k_gather_words(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, const uint32_t* __restrict__ map, uint32_t n, uint32_t words) {
    const size_t total = (size_t)n * words;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(i / words), c = (uint32_t)(i % words);
        dst[i] = src[(size_t)map[r] * words + c];
    }
}

GraphView Shard::view(uint32_t visible, int top) const {
    GraphView g{};
    g.vecs = vecs(); g.sqn = sqn(); g.links0 = links0();
    for (int l = 0; l < MAX_LEVELS; ++l) { g.linksU[l] = linksU(l); g.stride[l] = stride[l] ? stride[l] : 0xFFFFFFFFu; }
    g.deleted = deleted(); g.slot_key = slot_key(); g.n_slots = visible;
    g.dims = dims; g.dpad = dpad; g.chunks = dpad / 4; g.M = M; g.W0 = W0; g.metric = metric; g.top_level = top;
    return g;
}

int Shard::top_for(uint32_t n) const {
    int t = 0;
    for (int l = 1; l < n_levels; ++l) if (n > 0 && stride[l] <= n - 1) t = l;
    return t;
}

cudaEvent_t Shard::take_event() {   // pub_mu held
    if (!ev_pool.empty()) { cudaEvent_t e = ev_pool.back(); ev_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return e;
}

// The prefix of the graph a search may read: every insertion batch whose link step has finished on the GPU.
void Shard::snapshot(uint32_t& n, int& top) {
    std::lock_guard<std::mutex> g(pub_mu);
    while (!pubs.empty()) {
        const cudaError_t e = cudaEventQuery(pubs.front().ev);
        if (e != cudaSuccess) { if (e != cudaErrorNotReady) cudaGetLastError(); break; }
        vis_slots = pubs.front().n_slots; vis_top = pubs.front().top;
        ev_pool.push_back(pubs.front().ev); pubs.pop_front();
    }
    n = vis_slots; top = vis_top;
}

// usearch::Index::reserve (usearch.rs:99,206): every table grows; contents are preserved. With VMM-backed tables nothing moves
// (searches keep running); otherwise the tables are re-allocated under the exclusive table lock.
vs_status Shard::do_reserve(size_t cap) {
    if (cap <= capacity) return VS_OK;
    if (cap > max_slots) {
        if (capacity >= max_slots) return fail(VS_ERR_CAPACITY, "reserve: capacity exceeds what this GPU can hold (or 2^30 slots per shard)");
        cap = max_slots;
    }
    cudaSetDevice(dev.device);
    const double t0 = now_s();
    const bool in_place = a_vecs.vmm;
    std::unique_lock<std::shared_mutex> excl(tables, std::defer_lock);
    if (!in_place) { excl.lock(); for (int i = 0; i < ASYNC_EVENTS; ++i) cudaEventSynchronize(async_ev[i]); cudaStreamSynchronize(sstream); }
    const size_t old = capacity;
    cudaError_t e;
    size_t before = 0, after = 0;
    auto held = [&]() { size_t b = a_vecs.bytes + a_sqn.bytes + a_links0.bytes + a_slot_key.bytes + a_deleted.bytes + a_inbox_cnt.bytes + a_inbox_idx.bytes;
                        for (int l = 1; l < MAX_LEVELS; ++l) b += a_linksU[l].bytes; return b; };
    before = held();
#define GROW(arr, want, fill, keep) do { e = (arr).ensure((want), (fill), (keep), stream, nullptr); \
        if (e != cudaSuccess) { cudaGetLastError(); return fail(VS_ERR_CAPACITY, std::string("reserve: ") + cudaGetErrorString(e)); } } while (0)
    GROW(a_vecs, cap * dpad * 4, -1, old * dpad * 4);
    GROW(a_sqn, cap * 4, -1, old * 4);
    GROW(a_links0, cap * W0 * 4, 0xFF, old * W0 * 4);
    GROW(a_slot_key, cap * 8, 0xFF, old * 8);
    GROW(a_deleted, ((cap + 31) / 32) * 4, 0, ((old + 31) / 32) * 4);
    // level tables: level l exists while M^l < capacity
    int nl = 1; uint64_t s = M;
    while (nl < MAX_LEVELS && s < cap) { stride[nl] = (uint32_t)s; ++nl; s *= M; }
    stride[0] = 1;
    size_t total = cap; lt.row_base[0] = 0;
    for (int l = 1; l < nl; ++l) {
        const size_t nr = (cap + stride[l] - 1) / stride[l];
        if (!a_linksU[l].p && !a_linksU[l].bytes) a_linksU[l].init(dev.device, ((max_slots + stride[l] - 1) / stride[l]) * M * 4, use_vmm);
        GROW(a_linksU[l], nr * M * 4, 0xFF, rows[l] * M * 4);
        rows[l] = nr; lt.row_base[l] = (uint32_t)total; total += nr;
    }
    rows[0] = cap; n_levels = nl;
    // inbox index space covers the rows of every level; it is all-zero between batches, so the re-based offsets need no copy
    GROW(a_inbox_cnt, total * 4, 0, total_rows * 4);
    GROW(a_inbox_idx, total * 4, 0xFF, total_rows * 4);
    lt.inbox_cnt = reinterpret_cast<uint32_t*>(a_inbox_cnt.p); lt.inbox_idx = reinterpret_cast<uint32_t*>(a_inbox_idx.p);
#undef GROW
    CU(cudaStreamSynchronize(stream));
    total_rows = total;
    after = held();
    device_bytes += after - before;
    capacity = cap; cap_pub.store(cap);
    st_reserve_calls++; st_reserve_seconds += now_s() - t0;
    return VS_OK;
}

// Carves the per-batch scratch (candidate lists, intra-batch neighbours, edge/inbox arrays).
vs_status Shard::setup_build_scratch() {
    const uint32_t B = max_batch, efp = build_efp(efA), kbp = KNN_BATCH;
    efpA = efp;
    size_t bytes = 0; size_t off_ci[MAX_LEVELS], off_cd[MAX_LEVELS], off_cn[MAX_LEVELS], off_ki[MAX_LEVELS], off_kd[MAX_LEVELS], off_kis[2][MAX_LEVELS], off_kds[2][MAX_LEVELS];
    auto take = [&](size_t b) { size_t o = bytes; bytes += (b + 255) & ~(size_t)255; return o; };
    size_t items_total = 0; uint64_t s = 1;
    for (int l = 0; l < MAX_LEVELS; ++l) {
        const size_t it = B / s + 2; items_total += it;
        off_ci[l] = take(it * efp * 4); off_cd[l] = take(it * efp * 4); off_cn[l] = take(it * 4);
        off_ki[l] = take(it * kbp * 4); off_kd[l] = take(it * kbp * 4);
        for (int e = 0; e < 2; ++e) { off_kis[e][l] = take(it * kbp * 4); off_kds[e][l] = take(it * kbp * 4); }   // further sets: the neighbours of batches i+1, i+2 are computed while batch i searches and links
        s *= M; if (s > 0xFFFFFFFFull) s = 0xFFFFFFFFull;
    }
    edge_cap = items_total * M;
    const size_t ec = edge_cap;
    const size_t o_ne = take(16), o_er = take(ec * 4), o_es = take(ec * 4), o_tr = take(ec * 4), o_ts = take(ec * 4), o_tl = take(ec),
                 o_if = take(ec * 4), o_io = take((ec + 1) * 4), o_in = take(ec * 4);
    CU(build_scratch.ensure(bytes));
    unsigned char* p = build_scratch.p;
    CU(cudaMemsetAsync(p, 0, bytes, stream));
    for (int l = 0; l < MAX_LEVELS; ++l) {
        bb.cand_id[l] = (uint32_t*)(p + off_ci[l]); bb.cand_d[l] = (float*)(p + off_cd[l]); bb.cand_n[l] = (uint32_t*)(p + off_cn[l]);
        bb.knn_id[l] = (uint32_t*)(p + off_ki[l]); bb.knn_d[l] = (float*)(p + off_kd[l]);
        knn_set_id[0][l] = bb.knn_id[l]; knn_set_d[0][l] = bb.knn_d[l];
        for (int e = 0; e < 2; ++e) { knn_set_id[e + 1][l] = (uint32_t*)(p + off_kis[e][l]); knn_set_d[e + 1][l] = (float*)(p + off_kds[e][l]); }
    }
    lt.n_edges = (uint32_t*)(p + o_ne); lt.n_touched = lt.n_edges + 1; lt.inbox_cursor = lt.n_edges + 2;
    lt.edge_row = (uint32_t*)(p + o_er); lt.edge_src = (uint32_t*)(p + o_es);
    lt.touched_row = (uint32_t*)(p + o_tr); lt.touched_slot = (uint32_t*)(p + o_ts); lt.touched_level = (uint8_t*)(p + o_tl);
    lt.inbox_fill = (uint32_t*)(p + o_if); lt.inbox_off = (uint32_t*)(p + o_io); lt.inbox = (uint32_t*)(p + o_in);
    CU(exact_scratch.ensure(exact_scratch_bytes_upto(B, KNN_BATCH, dev.sm_count)));
    return VS_OK;
}

// Levels and item ranges of the batch [first, first + count).
void Shard::batch_shape(uint32_t first, uint32_t count, BuildBatch& b) const {
    b.first_slot = first; b.count = count;
    b.new_top = top_for(first + count);
    uint32_t off = 0;
    for (int l = 0; l < MAX_LEVELS; ++l) {
        b.item_off[l] = off;
        if (l <= b.new_top) {
            const uint32_t s = stride[l];
            const uint32_t fr = (first + s - 1) / s, lr = (first + count - 1) / s;
            b.first_row[l] = fr; b.items[l] = lr >= fr ? lr - fr + 1 : 0;
        } else { b.first_row[l] = 0; b.items[l] = 0; }
        off += b.items[l];
    }
    b.item_off[MAX_LEVELS] = off; b.total_items = off;
}

// Exact intra-batch neighbours (K1/K2/K1f) of the batch on stream ks, into knn buffer set `buf`. They depend on the batch's
// own rows only, not on the graph, so they may run any time after the rows are in place.
vs_status Shard::batch_knn(uint32_t first, uint32_t count, int buf, cudaStream_t ks) {
    BuildBatch b = bb;
    batch_shape(first, count, b);
    for (int l = 0; l <= b.new_top; ++l) {
        if (b.items[l] < 2) continue;
        uint32_t* kid = knn_set_id[buf][l]; float* kd = knn_set_d[buf][l];
        const uint32_t s = stride[l], fs = b.first_row[l] * s;
        ExactCorpus c{}; c.vecs = vecs() + (size_t)fs * dpad; c.sqn = sqn() + fs; c.row_stride = (size_t)s * dpad; c.row_step = s;
        c.first_slot = fs; c.n = b.items[l]; c.deleted = nullptr; c.slot_key = nullptr; c.n_slots = first + count;
        ExactQueries q{}; q.q = c.vecs; q.stride = c.row_stride; q.nq = b.items[l]; q.dims = dims; q.dpad = dpad; q.self_first = fs; q.self_step = s;
        const uint32_t k = std::min<uint32_t>(KNN_BATCH, b.items[l] - 1);
        ExactScratch sc{exact_scratch.p, exact_scratch.n};
        if (build_tc && exact_tc_supported(c, q, k)) {   // big level-0 blocks: tensor-core scan + exact re-rank (same results)
            size_t want = exact_tc_scratch_bytes(q.nq, c.n, k + 1, dpad, dev.sm_count);
            if (want > tc_scratch.n) {   // size it once for every batch shape up to max_batch: a cudaFree + cudaMalloc per growing batch cost up to 300 ms
                for (uint32_t bs = 4096; bs <= max_batch; bs += 128) want = std::max(want, exact_tc_scratch_bytes(bs, bs, KNN_BATCH + 1, dpad, dev.sm_count));
                want = std::max(want, exact_tc_scratch_bytes(max_batch, max_batch, KNN_BATCH + 1, dpad, dev.sm_count));
                CU(cudaStreamSynchronize(stream)); CU(cudaStreamSynchronize(ks));   // the scratch is about to move
            }
            CU(tc_scratch.ensure(want));
            ExactScratch ts{tc_scratch.p, tc_scratch.n};
            CU(launch_exact_tc(c, q, metric, k, KNN_BATCH, kid, nullptr, kd, ts, sc, dev, ks, &tc_build_stats));
            continue;
        }
        CU(launch_exact(c, q, metric, k, KNN_BATCH, kid, nullptr, kd, sc, dev, ks));
    }
    return VS_OK;
}

// K4a for slots [first, first + count): the batch's search on the frozen graph. `b` comes back filled for batch_link.
vs_status Shard::batch_search(uint32_t first, uint32_t count, int buf, BuildBatch& b) {
    b = bb;   // scratch pointers
    batch_shape(first, count, b);
    for (int l = 0; l < MAX_LEVELS; ++l) {
        if (l > b.new_top || b.items[l] < 2) { b.knn_id[l] = nullptr; b.knn_d[l] = nullptr; }
        else { b.knn_id[l] = knn_set_id[buf][l]; b.knn_d[l] = knn_set_d[buf][l]; }
    }
    auto mark = [&]() { if (build_trace) { cudaEvent_t e; if (cudaEventCreate(&e) == cudaSuccess) { cudaEventRecord(e, stream); trace_ev.push_back(e); } } };
    mark();
    GraphView frozen = view(first, b.new_top);
    CU(launch_build_search(frozen, b, efA, dev, stats_on.load() ? counters : nullptr, stream));
    mark();
    return VS_OK;
}

// K4b + K5: once the batch's exact neighbours are ready (`knn_ready`, null = already ordered on this stream), heuristic
// selection and reverse-link merge.
vs_status Shard::batch_link(const BuildBatch& b, cudaEvent_t knn_ready) {
    auto mark = [&]() { if (build_trace) { cudaEvent_t e; if (cudaEventCreate(&e) == cudaSuccess) { cudaEventRecord(e, stream); trace_ev.push_back(e); } } };
    if (knn_ready) CU(cudaStreamWaitEvent(stream, knn_ready, 0));
    mark();
    CU(cudaMemsetAsync(lt.n_edges, 0, 16, stream));
    GraphView linked = view(b.first_slot + b.count, b.new_top);
    CU(launch_link(linked, b, lt, efpA, KNN_BATCH, dev, stats_on.load() ? counters : nullptr, stream));
    mark();
    st_build_batches++;
    return VS_OK;
}

vs_status Shard::wait_async(cudaStream_t st) {
    if (async_used.load()) for (int i = 0; i < ASYNC_EVENTS; ++i) CU(cudaStreamWaitEvent(st, async_ev[i], 0));
    return VS_OK;
}

// Lands rows [base, base + n): per batch - host key map, tombstones of replaced keys, keys + rows into the tables, exact
// intra-batch neighbours (side stream), graph search (K4a), selection + reverse links (K4b / K5), publication.
// Replace = tombstone the old slot + a new slot under the same key (usearch.rs:214-225).
vs_status Shard::land(const uint64_t* keys, const float* rows_src, size_t n, RowSrc src) {
    if (n == 0) return VS_OK;
    cudaSetDevice(dev.device);
    // capacity growth: the reference's threshold / increment rule (usearch.rs:62-66,201-212)
    while (capacity - n_slots < std::max(reserve_increment / 3, n)) {
        const size_t want = capacity + std::max(reserve_increment, n);
        vs_status s = do_reserve(want); if (s != VS_OK) return s;
        if (capacity < want && capacity - n_slots < n) return fail(VS_ERR_CAPACITY, "add: the GPU cannot hold more vectors");
        if (capacity >= max_slots) break;
    }
    if (capacity - n_slots < n) return fail(VS_ERR_CAPACITY, "add: the GPU cannot hold more vectors");
    const uint32_t base = n_slots;
    CU(cudaEventRecord(ev0, stream));
    // bootstrap block (all-pairs), then batches of at most 1/batch_div of the landed graph so
    // that almost every candidate a new node needs is already linked (recall vs batch size: DESIGN.md)
    std::vector<std::pair<uint32_t, uint32_t>> bl;
    for (size_t done = 0; done < n;) {
        const uint32_t cur = base + (uint32_t)done;
        uint32_t bsz = cur == 0 ? 1024u : std::max(256u, cur / batch_div);
        bsz = (uint32_t)std::min<size_t>({(size_t)bsz, (size_t)max_batch, n - done});
        bl.emplace_back(cur, bsz); done += bsz;
    }
    std::vector<uint32_t> dead;
    // host half of a batch + its rows into the tables, on stream s
    auto stage_batch = [&](size_t i, cudaStream_t s) -> vs_status {
        const uint32_t first = bl[i].first, cnt = bl[i].second; const size_t off = first - base;
        if (src != RowSrc::IN_PLACE) {
            dead.clear();
            {
                std::lock_guard<std::mutex> g(map_mu);
                key_slot.reserve(key_slot.size() + cnt);
                for (uint32_t j = 0; j < cnt; ++j) {
                    const uint32_t old = key_slot.put(keys[off + j], first + j);
                    if (old != FlatMap::NO_SLOT) dead.push_back(old);
                }
                live += cnt - dead.size();
            }
            if (!dead.empty()) {
                CU(tomb.ensure(std::max<size_t>(dead.size(), max_batch)));
                CU(cudaMemcpyAsync(tomb.p, dead.data(), dead.size() * 4, cudaMemcpyHostToDevice, s));
                CU(launch_tombstone(deleted(), tomb.p, (uint32_t)dead.size(), s));
            }
            CU(cudaMemcpyAsync(slot_key() + first, keys + off, (size_t)cnt * 8, cudaMemcpyHostToDevice, s));
            if (src == RowSrc::HOST && dims == dpad) {     // rows go straight into the table; only the norms remain
                CU(cudaMemcpyAsync(vecs() + (size_t)first * dpad, rows_src + off * dims, (size_t)cnt * dims * 4, cudaMemcpyHostToDevice, s));
                CU(launch_row_prepare(nullptr, vecs(), sqn(), first, cnt, dims, dpad, s));
            } else if (src == RowSrc::HOST) {
                CU(stage.ensure((size_t)max_batch * dims * 4));
                CU(cudaMemcpyAsync(stage.p, rows_src + off * dims, (size_t)cnt * dims * 4, cudaMemcpyHostToDevice, s));
                CU(launch_row_prepare((const float*)stage.p, vecs(), sqn(), first, cnt, dims, dpad, s));
            } else {
                CU(launch_row_prepare(rows_src + off * dims, vecs(), sqn(), first, cnt, dims, dpad, s));
            }
        }
        return VS_OK;
    };
    auto publish = [&](size_t i) -> vs_status {
        std::lock_guard<std::mutex> g(pub_mu);
        cudaEvent_t e = take_event();
        if (!e) return fail(VS_ERR_CUDA, "flush: cannot create an event");
        CU(cudaEventRecord(e, stream));
        const uint32_t upto = bl[i].first + bl[i].second;
        pubs.push_back(Pub{e, upto, top_for(upto)});
        return VS_OK;
    };
    { vs_status w = wait_async(stream); if (w != VS_OK) return w; w = wait_async(side); if (w != VS_OK) return w; }   // searches launched asynchronously on user streams read the tables (and, in exact mode, scratch) - order after them
    vs_status st = VS_OK;
    if (!build_overlap) {            // one stream: rows, neighbours, search, link, batch after batch
        for (size_t i = 0; i < bl.size() && st == VS_OK; ++i) {
            BuildBatch b;
            if ((st = stage_batch(i, stream)) != VS_OK) break;
            if ((st = batch_knn(bl[i].first, bl[i].second, 0, stream)) != VS_OK) break;
            if ((st = batch_search(bl[i].first, bl[i].second, 0, b)) != VS_OK) break;
            if ((st = batch_link(b, nullptr)) != VS_OK) break;
            st = publish(i);
        }
    } else {
        // The rows of a batch and its exact intra-batch neighbours (K1/K2/K1f) do not depend on the graph, so they go to the side
        // stream - next to the graph search of the same batch, and build_pipeline = LA batches AHEAD: batch i+LA is staged and its
        // neighbours are computed while batch i searches and links (the persistent search kernel fills every SM, the link kernels
        // leave room). LA+1 buffer sets rotate; a set is reused once the link step that read it has finished.
        const int LA = std::max(1, build_pipeline), NS = LA + 1;
        bool linked_once[KNN_SETS] = {};
        CU(cudaEventRecord(rows_ready[0], stream));                 // order the side stream after whatever ran on the main stream
        CU(cudaStreamWaitEvent(side, rows_ready[0], 0));
        auto enqueue_side = [&](size_t i) -> vs_status {
            const int buf = (int)(i % NS);
            if (linked_once[buf]) CU(cudaStreamWaitEvent(side, link_done[buf], 0));
            vs_status s = stage_batch(i, side); if (s != VS_OK) return s;
            CU(cudaEventRecord(rows_ready[buf], side));
            s = batch_knn(bl[i].first, bl[i].second, buf, side); if (s != VS_OK) return s;
            CU(cudaEventRecord(knn_done[buf], side));
            return VS_OK;
        };
        // Host order matters as much as stream order: a batch costs ~1 ms of host time to enqueue against ~5 ms on the GPU, so the
        // graph search of batch i goes to the GPU FIRST and the host prepares the side-stream work while it runs.
        for (size_t i = 0; i < (size_t)LA && i < bl.size() && st == VS_OK; ++i) st = enqueue_side(i);
        for (size_t i = 0; i < bl.size() && st == VS_OK; ++i) {
            const int buf = (int)(i % NS);
            BuildBatch b;
            if (cudaStreamWaitEvent(stream, rows_ready[buf], 0) != cudaSuccess) { st = fail(VS_ERR_CUDA, "flush: cudaStreamWaitEvent"); break; }
            if ((st = batch_search(bl[i].first, bl[i].second, buf, b)) != VS_OK) break;
            if (i + LA < bl.size()) { if ((st = enqueue_side(i + LA)) != VS_OK) break; }
            if ((st = batch_link(b, knn_done[buf])) != VS_OK) break;
            if (cudaEventRecord(link_done[buf], stream) != cudaSuccess) { st = fail(VS_ERR_CUDA, "flush: cudaEventRecord"); break; }
            linked_once[buf] = true;
            st = publish(i);
        }
    }
    if (st == VS_OK && cudaEventRecord(ev1, stream) != cudaSuccess) st = fail(VS_ERR_CUDA, "flush: cudaEventRecord");
    const cudaError_t se = cudaStreamSynchronize(stream), se2 = cudaStreamSynchronize(side);
    if (st == VS_OK && (se != cudaSuccess || se2 != cudaSuccess)) st = fail(VS_ERR_CUDA, std::string("flush: ") + cudaGetErrorString(se != cudaSuccess ? se : se2));
    if (st != VS_OK) {   // keep the published prefix consistent with what is in the tables: nothing of this call becomes visible
        std::string keep = g_err;
        { std::lock_guard<std::mutex> g(pub_mu); for (auto& p : pubs) ev_pool.push_back(p.ev); pubs.clear(); }
        g_err = keep;
        return st;
    }
    float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1); st_build_seconds += ms * 1e-3;
    if (build_trace && !trace_ev.empty()) {   // per-phase device time and the gaps between batches
        double ph[3] = {0, 0, 0}, gaps = 0, worst = 0; size_t worst_i = 0;
        const size_t nb = trace_ev.size() / 4;
        for (size_t i = 0; i < nb; ++i) {
            float t;
            for (int k = 0; k < 3; ++k) { cudaEventElapsedTime(&t, trace_ev[4 * i + k], trace_ev[4 * i + k + 1]); ph[k] += t; }
            cudaEventElapsedTime(&t, trace_ev[4 * i], trace_ev[4 * i + 3]);
            if (t > worst) { worst = t; worst_i = i; }
            if (i + 1 < nb) { cudaEventElapsedTime(&t, trace_ev[4 * i + 3], trace_ev[4 * i + 4]); gaps += t; }
        }
        float w3[3] = {0, 0, 0};
        for (int k = 0; k < 3 && nb; ++k) cudaEventElapsedTime(&w3[k], trace_ev[4 * worst_i + k], trace_ev[4 * worst_i + k + 1]);
        fprintf(stderr, "[vsann build trace] %zu batches: total %.1f ms = search %.1f + batch-knn %.1f + link %.1f + gaps %.1f; slowest batch #%zu %.2f ms (search %.2f, knn wait %.2f, link %.2f)\n",
                nb, ms, ph[0], ph[1], ph[2], gaps, worst_i, worst, w3[0], w3[1], w3[2]);
        for (cudaEvent_t e : trace_ev) cudaEventDestroy(e);
        trace_ev.clear();
    }
    n_slots = base + (uint32_t)n;
    {   // everything has landed: publish the lot
        std::lock_guard<std::mutex> g(pub_mu);
        for (auto& p : pubs) ev_pool.push_back(p.ev);
        pubs.clear();
        vis_slots = n_slots; vis_top = top_for(n_slots);
    }
    st_build_vectors += n;
    return VS_OK;
}

// Lands every staged add (takes the staging buffers batch by batch, so adders are only ever blocked for a swap).
vs_status Shard::flush_locked() {
    for (;;) {
        std::vector<uint64_t> keys; std::vector<float> rows;
        {
            std::lock_guard<std::mutex> g(pend_mu);
            if (pend_keys.empty()) break;
            keys.swap(pend_keys); rows.swap(pend_vecs); pend_pos.clear();
        }
        flushing.store(true);
        const vs_status s = land(keys.data(), rows.data(), keys.size(), RowSrc::HOST);
        flushing.store(false);
        if (s != VS_OK) return s;
    }
    return maybe_compact_locked();
}

vs_status Shard::flush() { std::lock_guard<std::mutex> g(mu); return flush_locked(); }

// A search never waits for a running flush; if nobody is flushing and adds are staged, it lands them first (the single-threaded
// idiom "add, then search" sees its own adds, as the reference's tests expect after polling count, usearch.rs:352-358).
void Shard::maybe_flush_nonblocking() {
    { std::lock_guard<std::mutex> g(pend_mu); if (pend_keys.empty()) return; }
    std::unique_lock<std::mutex> l(mu, std::try_to_lock);
    if (l.owns_lock()) flush_locked();
}

size_t Shard::count() {
    std::lock_guard<std::mutex> g(mu);
    flush_locked();
    std::lock_guard<std::mutex> m(map_mu);
    return live;
}

vs_status Shard::reserve(size_t cap) { std::lock_guard<std::mutex> g(mu); return do_reserve(cap); }

vs_status Shard::add_host(size_t n, const uint64_t* keys, const float* rows) {
    if (n == 0) return VS_OK;
    if (n >= 1024) {   // bulk: the caller's rows go to the GPU batch by batch straight from its buffer (pinned or pageable), no staging copy
        std::lock_guard<std::mutex> g(mu);
        vs_status s = flush_locked(); if (s != VS_OK) return s;
        flushing.store(true);
        s = land(keys, rows, n, RowSrc::HOST);
        flushing.store(false);
        if (s != VS_OK) return s;
        return maybe_compact_locked();
    }
    bool full = false;
    {
        std::lock_guard<std::mutex> g(pend_mu);
        for (size_t i = 0; i < n; ++i) {
            const uint32_t at = pend_pos.get(keys[i]);
            if (at != FlatMap::NO_SLOT) {   // replaced before it landed: overwrite in place
                std::memcpy(pend_vecs.data() + (size_t)at * dims, rows + i * dims, dims * sizeof(float));
                continue;
            }
            pend_pos.put(keys[i], (uint32_t)pend_keys.size());
            pend_keys.push_back(keys[i]);
            pend_vecs.insert(pend_vecs.end(), rows + i * dims, rows + (i + 1) * dims);
        }
        full = pend_keys.size() >= max_batch;
    }
    if (full) {   // whoever fills the staging buffer lands it - unless a flush is already running (it will pick the rows up, or the next caller will)
        std::unique_lock<std::mutex> l(mu, std::try_to_lock);
        if (l.owns_lock()) return flush_locked();
    }
    return VS_OK;
}

vs_status Shard::add_device(size_t n, const uint64_t* keys, const float* d_rows) {
    if (n == 0) return VS_OK;
    std::lock_guard<std::mutex> g(mu);
    vs_status s = flush_locked(); if (s != VS_OK) return s;
    cudaSetDevice(dev.device);
    CU(cudaStreamSynchronize(nullptr));   // rows were produced on the caller's (legacy default) stream or earlier
    flushing.store(true);
    s = land(keys, d_rows, n, RowSrc::DEVICE);
    flushing.store(false);
    if (s != VS_OK) return s;
    return maybe_compact_locked();
}

// Tombstones queued by remove(): applied in one launch by whoever gets there first; the caller returns once ITS slot is marked.
vs_status Shard::apply_tombstones() {
    std::lock_guard<std::mutex> g(tomb_mu);
    std::vector<uint32_t> q;
    { std::lock_guard<std::mutex> l(tombq_mu); q.swap(tomb_queue); }
    if (q.empty()) return VS_OK;      // a caller ahead of us applied ours while it held tomb_mu
    cudaSetDevice(dev.device);
    CU(tomb1.ensure(std::max<size_t>(q.size(), 1024)));
    CU(cudaMemcpyAsync(tomb1.p, q.data(), q.size() * 4, cudaMemcpyHostToDevice, tstream));
    CU(launch_tombstone(deleted(), tomb1.p, (uint32_t)q.size(), tstream));
    CU(cudaStreamSynchronize(tstream));
    return VS_OK;
}

// usearch::Index::remove (usearch.rs:215,245): unmap the key, tombstone its slot. No flush, no insertion batch is cut short:
// a staged add of the key is simply dropped; a landed one gets its bit set by a 4-byte copy + one tiny launch on a stream of its own.
vs_status Shard::remove(uint64_t key) {
    bool was_pending = false;
    {
        std::lock_guard<std::mutex> g(pend_mu);
        const uint32_t at = pend_pos.erase(key);
        if (at != FlatMap::NO_SLOT) {
            was_pending = true;
            const size_t last = pend_keys.size() - 1;
            if (at != last) {
                pend_keys[at] = pend_keys[last];
                std::memcpy(pend_vecs.data() + (size_t)at * dims, pend_vecs.data() + last * dims, dims * sizeof(float));
                pend_pos.put(pend_keys[at], at);
            }
            pend_keys.pop_back(); pend_vecs.resize(last * dims);
        }
    }
    uint32_t slot = FlatMap::NO_SLOT;
    for (int attempt = 0; attempt < 2; ++attempt) {
        {
            std::lock_guard<std::mutex> g(map_mu);
            slot = key_slot.erase(key);
            if (slot != FlatMap::NO_SLOT) { --live; std::lock_guard<std::mutex> l(tombq_mu); tomb_queue.push_back(slot); }   // unmapped and queued in one step: compaction sees both or neither
        }
        if (slot != FlatMap::NO_SLOT || was_pending || attempt == 1 || !flushing.load()) break;
        std::lock_guard<std::mutex> w(mu);   // the key may belong to a batch a running flush has not registered yet: wait for it once
    }
    if (slot == FlatMap::NO_SLOT) return was_pending ? VS_OK : fail(VS_NOT_FOUND, "remove: key not found");
    vs_status s = apply_tombstones(); if (s != VS_OK) return s;
    std::unique_lock<std::mutex> l(mu, std::try_to_lock);
    if (l.owns_lock()) return maybe_compact_locked();
    return VS_OK;
}

// ---- compaction ----------------------------------------------------------------------------------------------------
// usearch frees the slot of a removed key and reuses it for the next add (usearch.rs:214-221,245). Here slots
// are positional (level = divisibility of the slot number, rows in arrival order), so instead of patching single slots the index is
// COMPACTED once tombstones exceed compact_frac of the slots: live rows are gathered to the front on the device (order kept) and the
// graph is rebuilt by the normal insertion pipeline - at ~2 M vectors/s that costs less per replaced key than one host-side hash
// probe per search would, bounds HBM use to live / (1 - compact_frac), and leaves no dead nodes for traversals to wade through.
vs_status Shard::maybe_compact_locked() {
    size_t lv;
    { std::lock_guard<std::mutex> m(map_mu); lv = live; }
    const size_t dead = n_slots > lv ? n_slots - lv : 0;
    if (n_slots < compact_min || (double)dead <= compact_frac * (double)n_slots) return VS_OK;
    return compact_locked();
}

vs_status Shard::compact_locked() {
    cudaSetDevice(dev.device);
    const double t0 = now_s();
    std::unique_lock<std::shared_mutex> excl(tables);      // no search may be in flight: slots are renumbered
    for (int i = 0; i < ASYNC_EVENTS; ++i) CU(cudaEventSynchronize(async_ev[i]));
    CU(cudaStreamSynchronize(sstream)); CU(cudaStreamSynchronize(tstream)); CU(cudaStreamSynchronize(stream)); CU(cudaStreamSynchronize(side));
    const uint32_t n_old = n_slots;
    std::vector<uint32_t> del((n_old + 31) / 32);
    std::vector<uint32_t> live_slots; live_slots.reserve(n_old);
    std::vector<uint32_t> remap(n_old, FlatMap::NO_SLOT);
    uint32_t n_live = 0;
    {
        std::lock_guard<std::mutex> m(map_mu);            // removes wait here; every tombstone they queued is applied before the bitmask is read
        vs_status ts = apply_tombstones(); if (ts != VS_OK) return ts;
        if (n_old) CU(cudaMemcpy(del.data(), deleted(), del.size() * 4, cudaMemcpyDeviceToHost));
        for (uint32_t s = 0; s < n_old; ++s) if (!((del[s >> 5] >> (s & 31)) & 1u)) { remap[s] = (uint32_t)live_slots.size(); live_slots.push_back(s); }
        n_live = (uint32_t)live_slots.size();
        if (key_slot.size() != n_live) return fail(VS_ERR_INVALID, "compact: key map and tombstones disagree");
        key_slot.update_each([&](uint64_t, uint32_t s) { return remap[s]; });
    }
    // gather live rows to the front, chunk by chunk through a scratch buffer (destination <= source, chunks in ascending order)
    if (n_live) {
        CU(cmap.ensure(n_live));
        CU(cudaMemcpyAsync(cmap.p, live_slots.data(), (size_t)n_live * 4, cudaMemcpyHostToDevice, stream));
        const uint32_t CH = 1u << 18;
        const size_t row_words = std::max<size_t>(dpad, 2);
        CU(stage.ensure((size_t)CH * row_words * 4));
        auto gather = [&](void* table, uint32_t words) -> vs_status {
            for (uint32_t c0 = 0; c0 < n_live; c0 += CH) {
                const uint32_t cnt = std::min(CH, n_live - c0);
                const size_t total = (size_t)cnt * words;
                const uint32_t blocks = (uint32_t)std::min<size_t>((total + 255) / 256, (size_t)dev.sm_count * 16);
                k_gather_words<<<blocks, 256, 0, stream>>>((const uint32_t*)table, (uint32_t*)stage.p, cmap.p + c0, cnt, words);
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync((uint32_t*)table + (size_t)c0 * words, stage.p, total * 4, cudaMemcpyDeviceToDevice, stream));
            }
            return VS_OK;
        };
        vs_status s;
        if ((s = gather(vecs(), dpad)) != VS_OK) return s;
        if ((s = gather(sqn(), 1)) != VS_OK) return s;
        if ((s = gather(slot_key(), 2)) != VS_OK) return s;
    }
    // empty graph over the same tables
    if (n_old) {
        CU(cudaMemsetAsync(links0(), 0xFF, (size_t)n_old * W0 * 4, stream));
        for (int l = 1; l < n_levels; ++l) CU(cudaMemsetAsync(linksU(l), 0xFF, ((size_t)(n_old + stride[l] - 1) / stride[l]) * M * 4, stream));
        CU(cudaMemsetAsync(deleted(), 0, del.size() * 4, stream));
    }
    CU(cudaStreamSynchronize(stream));
    n_slots = 0;
    { std::lock_guard<std::mutex> g(pub_mu); for (auto& p : pubs) ev_pool.push_back(p.ev); pubs.clear(); vis_slots = 0; vis_top = 0; }
    const uint64_t bv = st_build_vectors.load();
    vs_status s = land(nullptr, nullptr, n_live, RowSrc::IN_PLACE);
    st_build_vectors.store(bv);     // a rebuild is not ingest
    st_compactions++; st_compact_seconds += now_s() - t0;
    return s;
}

// ---- search --------------------------------------------------------------------------------------------------------
// device-accessible queries/outputs on stream st. ANN needs no scratch; exact mode uses the search-side scratch under search_mu.
vs_status Shard::search_device(size_t nq, const float* d_q, size_t k, size_t ef, int exact, uint64_t* d_keys, float* d_dist, cudaStream_t st, bool sync,
                               const OutFan* fan) {
    if (nq == 0) return VS_OK;
    if (fan && fan->n && exact) return fail(VS_ERR_UNSUPPORTED, "search: result fan-out is for the graph search (exact = 0) only");
    maybe_flush_nonblocking();
    cudaSetDevice(dev.device);
    std::shared_lock<std::shared_mutex> rd(tables);
    uint32_t vis; int top; snapshot(vis, top);
    unsigned long long* ctr = stats_on.load() ? counters : nullptr;
    auto finish = [&]() -> vs_status {
        if (sync) { CU(cudaStreamSynchronize(st)); return VS_OK; }
        const unsigned i = async_next.fetch_add(1) % ASYNC_EVENTS;
        CU(cudaEventRecord(async_ev[i], st)); async_used.store(true);
        return VS_OK;
    };
    if (exact) {
        std::lock_guard<std::mutex> sg(search_mu);
        if (exact_async_pending) CU(cudaStreamWaitEvent(st, exact_async_ev, 0));   // the scratch may still be in use by an earlier asynchronous exact search
        CU(qpad.ensure(nq * dpad));
        CU(s_exact_scratch.ensure(exact_scratch_bytes((uint32_t)nq, (uint32_t)k, dev.sm_count)));
        // pad the queries to dpad floats per row (aligned float4 loads in the tile kernel)
        CU(pad_queries(d_q, qpad.p, (uint32_t)nq, dims, dpad, st));
        ExactCorpus c{}; c.vecs = vecs(); c.sqn = sqn(); c.row_stride = dpad; c.row_step = 1; c.first_slot = 0; c.n = vis;
        c.deleted = deleted(); c.slot_key = slot_key(); c.n_slots = vis;
        ExactQueries q{}; q.q = qpad.p; q.stride = dpad; q.nq = (uint32_t)nq; q.dims = dims; q.dpad = dpad; q.self_first = 0; q.self_step = 0;
        ExactScratch sc{s_exact_scratch.p, s_exact_scratch.n};
        // K1s, the HBM-streaming scan: always for <= 4 queries (measured on B200, 5M x 1536 / 12.5M x 96 / 1.25M x 768: 4.4 / 0.9 / 0.7 ms
        // against 5.3 / 1.3 / 0.9 ms for the tensor path, whose fixed costs - sampling pass, re-rank, select - only pay off
        // from ~8 queries), up to 16 queries when the tensor path does not apply (k > 128, small corpora)
        const bool tc_ok = exact_tc_supported(c, q, (uint32_t)k);
        const bool streamk = exact_stream_supported(c, q, (uint32_t)k, dev) && (nq <= 4 || !tc_ok);
        if (streamk) st_stream_queries += nq;
        if (!streamk && tc_ok) {
            // tcgen05 TF32 scan + exact FP32 re-rank (exact_tc.cu); same ids and distance bits as the FP32 kernel
            CU(s_tc_scratch.ensure(exact_tc_scratch_bytes((uint32_t)nq, c.n, (uint32_t)k, dpad, dev.sm_count)));
            ExactScratch ts{s_tc_scratch.p, s_tc_scratch.n};
            CU(launch_exact_tc(c, q, metric, (uint32_t)k, (uint32_t)k, nullptr, d_keys, d_dist, ts, sc, dev, st, &tc_stats));
        } else {
            CU(launch_exact(c, q, metric, (uint32_t)k, (uint32_t)k, nullptr, d_keys, d_dist, sc, dev, st));
        }
        if (!sync) { CU(cudaEventRecord(exact_async_ev, st)); exact_async_pending = true; }
        return finish();
    }
    // Tombstoned nodes are traversed but never returned (usearch.rs:245): widen the beam by the dead fraction so that they do not
    // crowd live results out of the list (compaction bounds the fraction by compact_frac).
    size_t lv; { std::lock_guard<std::mutex> m(map_mu); lv = live; }
    uint32_t efe = (uint32_t)std::max(ef ? ef : (size_t)efS, k);
    if (vis > lv && lv > 0) efe = (uint32_t)std::min<uint64_t>(1024, ((uint64_t)efe * vis + lv - 1) / lv);
    GraphView g = view(vis, top);
    cudaError_t e = launch_search(g, d_q, (uint32_t)nq, (uint32_t)k, efe, dev, d_keys, d_dist, ctr, st, fan);
    if (e == cudaErrorInvalidConfiguration) return fail(VS_ERR_UNSUPPORTED, "search: ef / dims too large for one warp's shared memory");
    if (e == cudaErrorNotSupported) { cudaGetLastError(); return fail(VS_ERR_UNSUPPORTED, "search: result fan-out needs the staged traversal kernels (rows too long for the staging area)"); }
    CU(e);
    return finish();
}

vs_status Shard::search_host(size_t nq, const float* qs, size_t k, size_t ef, int exact, uint64_t* keys_out, float* dist_out, size_t* found) {
    if (nq == 0) return VS_OK;
    maybe_flush_nonblocking();
    cudaSetDevice(dev.device);
    std::shared_lock<std::shared_mutex> rd(tables);  // lock order everywhere: tables, then search_mu
    std::lock_guard<std::mutex> sg(search_mu);      // one host-buffer search at a time per shard: they share the staging buffers
    CU(out_k.ensure(nq * k)); CU(out_d.ensure(nq * k));
    // Pinned (page-locked, UVA-mapped) query buffers are read by the traversal kernel in place: every warp
    // fetches its 4 d bytes over PCIe while the others compute, so the H2D copy leaves the critical path.
    // Pageable buffers, and the exact path (which streams the queries many times through TMA), are staged.
    const float* d_q = nullptr;
    if (!exact && zero_copy) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, qs) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) d_q = (const float*)at.devicePointer;
        else cudaGetLastError();
    }
    if (!d_q) {
        CU(qdense.ensure(nq * dims));
        CU(cudaMemcpyAsync(qdense.p, qs, nq * dims * 4, cudaMemcpyHostToDevice, sstream));
        d_q = qdense.p;
    }
    st_search_launches++; st_search_calls += nq;
    vs_status s;
    {   // search_device takes search_mu itself in exact mode: run its body inline here instead
        uint32_t vis; int top; snapshot(vis, top);
        unsigned long long* ctr = stats_on.load() ? counters : nullptr;
        if (exact) {
            if (exact_async_pending) { CU(cudaStreamWaitEvent(sstream, exact_async_ev, 0)); }
            CU(qpad.ensure(nq * dpad));
            CU(s_exact_scratch.ensure(exact_scratch_bytes((uint32_t)nq, (uint32_t)k, dev.sm_count)));
            CU(pad_queries(d_q, qpad.p, (uint32_t)nq, dims, dpad, sstream));
            ExactCorpus c{}; c.vecs = vecs(); c.sqn = sqn(); c.row_stride = dpad; c.row_step = 1; c.first_slot = 0; c.n = vis;
            c.deleted = deleted(); c.slot_key = slot_key(); c.n_slots = vis;
            ExactQueries q{}; q.q = qpad.p; q.stride = dpad; q.nq = (uint32_t)nq; q.dims = dims; q.dpad = dpad; q.self_first = 0; q.self_step = 0;
            ExactScratch sc{s_exact_scratch.p, s_exact_scratch.n};
            const bool tc_ok = exact_tc_supported(c, q, (uint32_t)k);
            const bool streamk = exact_stream_supported(c, q, (uint32_t)k, dev) && (nq <= 4 || !tc_ok);
            if (streamk) st_stream_queries += nq;
            if (!streamk && tc_ok) {
                CU(s_tc_scratch.ensure(exact_tc_scratch_bytes((uint32_t)nq, c.n, (uint32_t)k, dpad, dev.sm_count)));
                ExactScratch ts{s_tc_scratch.p, s_tc_scratch.n};
                CU(launch_exact_tc(c, q, metric, (uint32_t)k, (uint32_t)k, nullptr, out_k.p, out_d.p, ts, sc, dev, sstream, &tc_stats));
            } else {
                CU(launch_exact(c, q, metric, (uint32_t)k, (uint32_t)k, nullptr, out_k.p, out_d.p, sc, dev, sstream));
            }
        } else {
            size_t lv; { std::lock_guard<std::mutex> m(map_mu); lv = live; }
            uint32_t efe = (uint32_t)std::max(ef ? ef : (size_t)efS, k);
            if (vis > lv && lv > 0) efe = (uint32_t)std::min<uint64_t>(1024, ((uint64_t)efe * vis + lv - 1) / lv);
            GraphView g = view(vis, top);
            cudaError_t e = launch_search(g, d_q, (uint32_t)nq, (uint32_t)k, efe, dev, out_k.p, out_d.p, ctr, sstream);
            if (e == cudaErrorInvalidConfiguration) return fail(VS_ERR_UNSUPPORTED, "search: ef / dims too large for one warp's shared memory");
            CU(e);
        }
        // (results: writing them straight into pinned host memory from the kernel was measured and is no faster than
        // this 12 k B per query copy, so the explicit copy stays)
        CU(cudaMemcpyAsync(keys_out, out_k.p, nq * k * 8, cudaMemcpyDeviceToHost, sstream));
        CU(cudaMemcpyAsync(dist_out, out_d.p, nq * k * 4, cudaMemcpyDeviceToHost, sstream));
        if (found) {   // counted on the device, 4 bytes per query come back
            CU(found_d.ensure(nq));
            CU(found_pin.ensure(std::max<size_t>(nq, 16384)));
            CU(launch_count_found(out_k.p, (uint32_t)nq, (uint32_t)k, found_d.p, sstream));
            CU(cudaMemcpyAsync(found_pin.p, found_d.p, nq * 4, cudaMemcpyDeviceToHost, sstream));
        }
        CU(cudaStreamSynchronize(sstream));
        s = VS_OK;
    }
    if (found) for (size_t i = 0; i < nq; ++i) found[i] = found_pin.p[i];
    return s;
}

// ---- lifecycle -----------------------------------------------------------------------------------------------------
vs_status Shard::create(const ShardOptions& o, Shard** out) {
    *out = nullptr;
    cudaDeviceProp prop{};
    CU(cudaGetDeviceProperties(&prop, o.device));
    if (prop.major < 10) return fail(VS_ERR_CUDA, "create: device is not Blackwell (sm_100a kernels only)");
    CU(cudaSetDevice(o.device));
    Shard* sh = new Shard();
    sh->dims = o.dims; sh->dpad = (o.dims + 3u) & ~3u; sh->metric = o.metric;
    sh->M = o.M; sh->W0 = 2 * o.M; sh->efA = o.efA; sh->efS = o.efS;
    sh->max_batch = o.max_batch ? o.max_batch : 32768;
    if (const char* e = getenv("VSANN_MAX_BATCH")) { const long v = atol(e); if (v > 0) sh->max_batch = (uint32_t)v; }
    if (sh->max_batch < 32) sh->max_batch = 32;
    if (const char* e = getenv("VSANN_BATCH_DIV")) { const long v = atol(e); if (v > 0) sh->batch_div = (uint32_t)v; }
    sh->reserve_increment = o.capacity_hint ? o.capacity_hint : 1000000;   // usearch.rs:62
    sh->dev.device = o.device; sh->dev.sm_count = prop.multiProcessorCount; sh->dev.smem_per_block = prop.sharedMemPerBlockOptin;
    if (const char* e = getenv("VSANN_HASH_FACTOR")) { const long v = atol(e); if (v > 0) sh->dev.hash_factor = (uint32_t)v; }
    if (const char* e = getenv("VSANN_HASH_MIN")) { const long v = atol(e); if (v >= 64) sh->dev.hash_min = (uint32_t)v; }
    if (const char* e = getenv("VSANN_SMEM_LIST")) sh->dev.force_smem_list = atoi(e) != 0;
    if (const char* e = getenv("VSANN_STAGE_BYTES")) sh->dev.stage_bytes = (uint32_t)atol(e);
    if (const char* e = getenv("VSANN_GATHER4")) sh->dev.gather4 = atoi(e) != 0;
    if (const char* e = getenv("VSANN_SELECT_SMEM")) sh->dev.select_smem = atoi(e) != 0;
    if (const char* e = getenv("VSANN_ZERO_COPY")) sh->zero_copy = atoi(e) != 0;   // 0: always stage host queries with a copy
    if (const char* e = getenv("VSANN_BUILD_PIPELINE")) sh->build_pipeline = std::max(1, std::min(KNN_SETS - 1, atoi(e)));
    if (const char* e = getenv("VSANN_BUILD_OVERLAP")) sh->build_overlap = atoi(e) != 0;   // 0: batch neighbours before the graph search, same stream
    if (const char* e = getenv("VSANN_BUILD_TRACE")) sh->build_trace = atoi(e) != 0;   // per-phase build timing on stderr
    if (const char* e = getenv("VSANN_BUILD_TC")) sh->build_tc = atoi(e) != 0;   // 0: FP32 tile kernel for the batch neighbours (A/B, profiling)
    if (const char* e = getenv("VSANN_COMPACT_FRAC")) { const double v = atof(e); if (v > 0 && v < 1) sh->compact_frac = v; else if (v >= 1) sh->compact_frac = 2.0; }   // >= 1: never compact
    if (const char* e = getenv("VSANN_VMM")) sh->use_vmm = atoi(e) != 0;
    // the largest index this GPU could ever hold bounds the virtual ranges (slots < 2^30: two flag bits live in the beam's list ids)
    sh->max_slots = std::min<size_t>(0x3FFFFFF0ull, prop.totalGlobalMem / ((size_t)sh->dpad * 4));
    if (sh->max_slots < sh->reserve_increment) sh->max_slots = sh->reserve_increment;
    auto body = [&]() -> vs_status {
        CU(cudaStreamCreateWithFlags(&sh->stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&sh->side, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&sh->sstream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&sh->tstream, cudaStreamNonBlocking));
        for (int e = 0; e < KNN_SETS; ++e) {
            CU(cudaEventCreateWithFlags(&sh->knn_done[e], cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&sh->link_done[e], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&sh->rows_ready[e], cudaEventDisableTiming));
        }
        for (int e = 0; e < ASYNC_EVENTS; ++e) CU(cudaEventCreateWithFlags(&sh->async_ev[e], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sh->exact_async_ev, cudaEventDisableTiming));
        CU(cudaEventCreate(&sh->ev0)); CU(cudaEventCreate(&sh->ev1));
        CU(cudaMalloc(&sh->counters, 8 * sizeof(unsigned long long)));
        CU(cudaMemsetAsync(sh->counters, 0, 8 * sizeof(unsigned long long), sh->stream));
        CU(cudaMalloc(&sh->dev.work_ring, WORK_RING_SLOTS * sizeof(uint32_t)));
        sh->tc_stats.dev = sh->counters + 5;
        const size_t ms = sh->max_slots; const bool v = sh->use_vmm;
        sh->a_vecs.init(o.device, ms * sh->dpad * 4, v); sh->a_sqn.init(o.device, ms * 4, v); sh->a_links0.init(o.device, ms * sh->W0 * 4, v);
        sh->a_slot_key.init(o.device, ms * 8, v); sh->a_deleted.init(o.device, ((ms + 31) / 32) * 4 + 64, v);
        size_t tot = ms; { uint64_t s = sh->M; while (s < ms) { tot += (ms + s - 1) / s; s *= sh->M; } }
        sh->a_inbox_cnt.init(o.device, (tot + MAX_LEVELS) * 4, v); sh->a_inbox_idx.init(o.device, (tot + MAX_LEVELS) * 4, v);
        vs_status r = sh->do_reserve(sh->reserve_increment); if (r != VS_OK) return r;
        r = sh->setup_build_scratch(); if (r != VS_OK) return r;
        CU(cudaStreamSynchronize(sh->stream));
        return VS_OK;
    };
    const vs_status s = body();
    if (s != VS_OK) { std::string keep = g_err; sh->destroy(); g_err = keep; return s; }
    *out = sh; return VS_OK;
}

void Shard::destroy() {
    cudaSetDevice(dev.device);
    for (cudaStream_t s : {stream, side, sstream, tstream}) if (s) cudaStreamSynchronize(s);
    a_vecs.release(); a_sqn.release(); a_links0.release(); a_slot_key.release(); a_deleted.release(); a_inbox_cnt.release(); a_inbox_idx.release();
    for (int l = 0; l < MAX_LEVELS; ++l) a_linksU[l].release();
    cudaFree(counters); cudaFree(dev.work_ring);
    build_scratch.release(); exact_scratch.release(); tc_scratch.release(); stage.release(); tomb.release(); cmap.release(); tomb1.release();
    s_exact_scratch.release(); s_tc_scratch.release(); qpad.release(); qdense.release(); out_d.release(); out_k.release(); found_d.release(); found_pin.release();
    if (ev0) cudaEventDestroy(ev0); if (ev1) cudaEventDestroy(ev1); if (exact_async_ev) cudaEventDestroy(exact_async_ev);
    for (int e = 0; e < KNN_SETS; ++e) { if (knn_done[e]) cudaEventDestroy(knn_done[e]); if (link_done[e]) cudaEventDestroy(link_done[e]); if (rows_ready[e]) cudaEventDestroy(rows_ready[e]); }
    for (int e = 0; e < ASYNC_EVENTS; ++e) if (async_ev[e]) cudaEventDestroy(async_ev[e]);
    for (auto& p : pubs) cudaEventDestroy(p.ev);
    for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
    for (cudaStream_t s : {stream, side, sstream, tstream}) if (s) cudaStreamDestroy(s);
    cudaGetLastError();
    delete this;
}

vs_status Shard::get_stats(vs_stats* o) {
    std::memset(o, 0, sizeof(*o));
    unsigned long long c[8];
    cudaSetDevice(dev.device);
    {
        std::lock_guard<std::mutex> g(search_mu);
        CU(cudaStreamSynchronize(sstream));
        CU(cudaMemcpy(c, counters, sizeof(c), cudaMemcpyDeviceToHost));
    }
    { std::lock_guard<std::mutex> m(map_mu); o->count = live; }
    { std::lock_guard<std::mutex> p(pend_mu); o->pending = pend_keys.size(); }
    uint32_t vis; int top; snapshot(vis, top);
    o->slots = vis; o->capacity = capacity;
    o->search_queries = c[0]; o->search_dist_evals = c[1]; o->search_expansions = c[2];
    o->build_vectors = st_build_vectors; o->build_dist_evals = c[3]; o->build_batches = st_build_batches; o->build_expansions = c[4];
    o->build_seconds = st_build_seconds; o->device_bytes = device_bytes + build_scratch.n + exact_scratch.n + tc_scratch.n;
    o->top_level = (uint32_t)top; o->dims = dims;
    o->exact_tc_queries = tc_stats.queries; o->exact_tc_fallback = c[6]; o->exact_tc_second_scan = c[5];
    o->exact_tc_scans = tc_stats.scans; o->exact_tc_scan_seconds = tc_stats.scan_ms * 1e-3;
    o->exact_stream_queries = st_stream_queries; o->search_host_queries = st_search_calls; o->search_host_launches = st_search_launches;
    o->reserve_calls = st_reserve_calls; o->reserve_seconds = st_reserve_seconds; o->compactions = st_compactions; o->compact_seconds = st_compact_seconds;
    o->tables_grow_in_place = a_vecs.vmm ? 1 : 0;
    return VS_OK;
}

vs_status Shard::reset_stats() {
    cudaSetDevice(dev.device);
    std::lock_guard<std::mutex> g(search_mu);
    CU(cudaMemsetAsync(counters, 0, 8 * sizeof(unsigned long long), sstream));
    CU(cudaStreamSynchronize(sstream));
    st_build_vectors = 0; st_build_batches = 0; st_build_seconds = 0; st_stream_queries = 0; st_search_calls = 0; st_search_launches = 0;
    { const bool t = tc_stats.time_scan; tc_stats = TcStats(); tc_stats.time_scan = t; tc_stats.dev = counters + 5; }
    return VS_OK;
}

vs_status Shard::export_level(uint32_t level, uint32_t* links_out, size_t* n_rows, size_t* row_width, uint64_t* slot_keys_out, size_t* slots) {
    std::lock_guard<std::mutex> g(mu);
    vs_status s = flush_locked(); if (s != VS_OK) return s;
    if ((int)level >= n_levels) return fail(VS_ERR_INVALID, "export: no such level");
    const size_t st = stride[level], nr = n_slots ? (n_slots - 1) / st + 1 : 0, w = level == 0 ? W0 : M;
    *n_rows = nr; *row_width = w; if (slots) *slots = n_slots;
    cudaSetDevice(dev.device);
    if (links_out && nr) CU(cudaMemcpy(links_out, level == 0 ? links0() : linksU(level), nr * w * 4, cudaMemcpyDeviceToHost));
    if (slot_keys_out && n_slots) CU(cudaMemcpy(slot_keys_out, slot_key(), (size_t)n_slots * 8, cudaMemcpyDeviceToHost));
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

vs_status Shard::save(const char* path) {
    std::lock_guard<std::mutex> g(mu);
    vs_status s = flush_locked(); if (s != VS_OK) return s;
    cudaSetDevice(dev.device);
    CU(cudaStreamSynchronize(stream)); CU(cudaStreamSynchronize(tstream));
    FILE* f = fopen(path, "wb");
    if (!f) return fail(VS_ERR_INVALID, std::string("save: cannot open ") + path);
    FileCloser fc{f};
    FileHeader h{};
    std::memcpy(h.magic, "VSANNIDX", 8);
    h.version = FILE_VERSION; h.dims = dims; h.dpad = dpad; h.metric = (uint32_t)metric; h.M = M; h.efA = efA; h.efS = efS;
    h.n_levels_used = (uint32_t)top_for(n_slots) + 1; h.n_slots = n_slots; { std::lock_guard<std::mutex> m(map_mu); h.live = live; } h.max_batch = max_batch;
    Summer sum; sum.add(&h, sizeof h);
    if (fwrite(&h, sizeof h, 1, f) != 1) return fail(VS_ERR_INVALID, "save: short write");
    std::vector<unsigned char> buf(64u << 20);
    const size_t n = n_slots;
    if ((s = dump(f, sum, vecs(), n * dpad * 4, buf)) != VS_OK) return s;
    if ((s = dump(f, sum, sqn(), n * 4, buf)) != VS_OK) return s;
    if ((s = dump(f, sum, links0(), n * W0 * 4, buf)) != VS_OK) return s;
    for (uint32_t l = 1; l < h.n_levels_used; ++l) {
        const size_t nr = (n + stride[l] - 1) / stride[l];
        if ((s = dump(f, sum, linksU(l), nr * M * 4, buf)) != VS_OK) return s;
    }
    if ((s = dump(f, sum, slot_key(), n * 8, buf)) != VS_OK) return s;
    if ((s = dump(f, sum, deleted(), ((n + 31) / 32) * 4, buf)) != VS_OK) return s;
    if (fwrite(&sum.h, 8, 1, f) != 1 || fflush(f) != 0) return fail(VS_ERR_INVALID, "save: short write");
    return VS_OK;
}

vs_status Shard::load(const char* path, const vs_options* ov, Shard** out) {
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(VS_ERR_INVALID, std::string("load: cannot open ") + path);
    FileCloser fc{f};
    FileHeader h{};
    if (fread(&h, sizeof h, 1, f) != 1 || std::memcmp(h.magic, "VSANNIDX", 8) != 0) return fail(VS_ERR_INVALID, "load: not a vsann index file");
    if (h.version > FILE_VERSION) return fail(VS_ERR_UNSUPPORTED, "load: file format " + std::to_string(h.version) + " is newer than this library");
    if (h.version == 0 || h.dims == 0 || h.dpad != ((h.dims + 3u) & ~3u) || h.metric > 2 || h.M < 2 || h.M > 16 || h.n_slots >= 0x3FFFFFF0ull || h.live > h.n_slots)
        return fail(VS_ERR_INVALID, "load: corrupt header");
    ShardOptions o;
    o.dims = h.dims; o.metric = (int)h.metric; o.M = h.M; o.efA = h.efA; o.efS = h.efS;
    o.device = ov ? ov->device : 0; o.max_batch = ov && ov->max_batch ? (uint32_t)ov->max_batch : h.max_batch;
    const size_t hint = ov && ov->capacity_hint ? ov->capacity_hint : 1000000;
    o.capacity_hint = hint;
    Shard* sh = nullptr;
    vs_status s = Shard::create(o, &sh); if (s != VS_OK) return s;
    auto body = [&]() -> vs_status {
        std::lock_guard<std::mutex> g(sh->mu);
        const size_t n = (size_t)h.n_slots;
        vs_status r = sh->do_reserve(n + std::max<size_t>(hint / 3, 1)); if (r != VS_OK) return r;
        if ((uint32_t)sh->top_for((uint32_t)n) + 1 != h.n_levels_used) return fail(VS_ERR_INVALID, "load: level count does not match the slot count");
        Summer sum; sum.add(&h, sizeof h);
        std::vector<unsigned char> buf(64u << 20);
        if ((r = restore(f, sum, sh->vecs(), n * sh->dpad * 4, buf)) != VS_OK) return r;
        if ((r = restore(f, sum, sh->sqn(), n * 4, buf)) != VS_OK) return r;
        if ((r = restore(f, sum, sh->links0(), n * sh->W0 * 4, buf)) != VS_OK) return r;
        for (uint32_t l = 1; l < h.n_levels_used; ++l) {
            const size_t nr = (n + sh->stride[l] - 1) / sh->stride[l];
            if ((r = restore(f, sum, sh->linksU(l), nr * sh->M * 4, buf)) != VS_OK) return r;
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
        if (n) { CU(cudaMemcpy(sh->slot_key(), keys.data(), n * 8, cudaMemcpyHostToDevice)); CU(cudaMemcpy(sh->deleted(), del.data(), del.size() * 4, cudaMemcpyHostToDevice)); }
        size_t lv = 0;
        sh->key_slot.reserve(h.live);
        for (size_t sl = 0; sl < n; ++sl) {
            if (del[sl >> 5] >> (sl & 31) & 1u) continue;
            if (sh->key_slot.put(keys[sl], (uint32_t)sl) != FlatMap::NO_SLOT) return fail(VS_ERR_INVALID, "load: two live slots share a key");
            ++lv;
        }
        if (lv != h.live) return fail(VS_ERR_INVALID, "load: live count does not match the tombstones");
        sh->n_slots = (uint32_t)n; sh->live = lv; sh->vis_slots = sh->n_slots; sh->vis_top = sh->top_for(sh->n_slots);
        return VS_OK;
    };
    s = body();
    if (s != VS_OK) { std::string keep = g_err; sh->destroy(); g_err = keep; return s; }
    *out = sh; return VS_OK;
}

}  // namespace vsann
