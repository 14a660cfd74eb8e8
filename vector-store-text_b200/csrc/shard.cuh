// shard.cuh — one GPU's part of an index: tables in HBM, the insertion pipeline, searches, removes, compaction.
//
// What the reference's actor does around usearch (src/index/usearch.rs), per device:
//   new + reserve            usearch.rs:82-99        Shard::create
//   add / replace            usearch.rs:174-233      Shard::add_host / add_device -> staged, landed in batches (K4 / K5)
//   remove (tombstone)       usearch.rs:235-249      Shard::remove
//   search -> Matches        usearch.rs:251-306      Shard::search_*
//   size / capacity/reserve  usearch.rs:201-212,308-311
//
// Concurrency (the reference holds usearch under Arc<RwLock<_>>: READ lock for add, remove, search and size together,
// WRITE lock for reserve only - usearch.rs:98,201-221,245,276,309 - with up to 2 x T calls in flight, :118-131):
//   mu         graph mutations: one insertion pipeline (flush) or one compaction at a time
//   pend_mu    the staging buffers of vs_add (held for a memcpy)
//   map_mu     key -> slot map and the live count (held per batch of keys, never across GPU work)
//   tables     shared: a search is in flight; exclusive: compaction, or table growth when the tables cannot grow in place
//   search_mu  the search-side scratch buffers (host-buffer searches and exact-mode scratch)
//   pub_mu     which slots searches may see: an insertion batch becomes visible when its link step has finished on the GPU
// A search never waits for a running flush: it reads the published prefix of the graph, on its own stream.
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

#include "../../include/vsann.h"
#include "flatmap.h"
#include "kernels.h"
#include "vmm.h"

namespace vsann {

extern thread_local std::string g_err;
vs_status fail(vs_status s, const std::string& m);
#define CU(expr)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) return vsann::fail(VS_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
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

template <class T> struct PinBuf {
    T* p = nullptr; size_t n = 0;
    cudaError_t ensure(size_t want) {
        if (want <= n) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; n = 0;
        cudaError_t e = cudaHostAlloc((void**)&p, want * sizeof(T), cudaHostAllocDefault);
        if (e == cudaSuccess) n = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; n = 0; }
};

struct ShardOptions {
    uint32_t dims = 0; int metric = 0; uint32_t M = 16, efA = 128, efS = 64; size_t capacity_hint = 1000000; int device = 0; uint32_t max_batch = 32768;
};

enum class RowSrc { HOST, DEVICE, IN_PLACE };

struct Shard {
    // ---- options -------------------------------------------------------------------------------------------
    uint32_t dims = 0, dpad = 0; int metric = 0; uint32_t M = 16, W0 = 32, efA = 128, efS = 64;
    uint32_t max_batch = 32768, batch_div = 16; size_t reserve_increment = 1000000;
    DeviceInfo dev;
    bool build_overlap = true, build_tc = true, zero_copy = true, build_trace = false, use_vmm = true; int build_pipeline = 1;
    double compact_frac = 0.25; size_t compact_min = 4096;
    // ---- streams / events ----------------------------------------------------------------------------------
    cudaStream_t stream = nullptr, side = nullptr, sstream = nullptr, tstream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    static constexpr int KNN_SETS = 3;
    cudaEvent_t knn_done[KNN_SETS] = {}, link_done[KNN_SETS] = {}, rows_ready[KNN_SETS] = {};
    static constexpr int ASYNC_EVENTS = 16;
    cudaEvent_t async_ev[ASYNC_EVENTS] = {}; std::atomic<unsigned> async_next{0}; std::atomic<bool> async_used{false};
    cudaEvent_t exact_async_ev = nullptr; bool exact_async_pending = false;     // exact-mode scratch in use by an asynchronous search (under search_mu)
    // ---- locks ---------------------------------------------------------------------------------------------
    std::mutex mu, pend_mu, map_mu, search_mu, pub_mu, tomb_mu, tombq_mu;
    std::shared_mutex tables;
    std::atomic<bool> flushing{false};
    // ---- tables --------------------------------------------------------------------------------------------
    size_t capacity = 0, max_slots = 0; uint32_t n_slots = 0; int n_levels = 1; std::atomic<size_t> cap_pub{0};
    VArr a_vecs, a_sqn, a_links0, a_linksU[MAX_LEVELS], a_slot_key, a_deleted, a_inbox_cnt, a_inbox_idx;
    uint32_t stride[MAX_LEVELS] = {}; size_t rows[MAX_LEVELS] = {};
    LinkTables lt{}; size_t total_rows = 0, edge_cap = 0;
    float* vecs() const { return reinterpret_cast<float*>(a_vecs.p); }
    float* sqn() const { return reinterpret_cast<float*>(a_sqn.p); }
    uint32_t* links0() const { return reinterpret_cast<uint32_t*>(a_links0.p); }
    uint32_t* linksU(int l) const { return reinterpret_cast<uint32_t*>(a_linksU[l].p); }
    uint64_t* slot_key() const { return reinterpret_cast<uint64_t*>(a_slot_key.p); }
    uint32_t* deleted() const { return reinterpret_cast<uint32_t*>(a_deleted.p); }
    // ---- published state (pub_mu) --------------------------------------------------------------------------
    struct Pub { cudaEvent_t ev; uint32_t n_slots; int top; };
    std::deque<Pub> pubs; std::vector<cudaEvent_t> ev_pool; uint32_t vis_slots = 0; int vis_top = 0;
    // ---- host maps (map_mu) --------------------------------------------------------------------------------
    FlatMap key_slot; size_t live = 0;
    // ---- staging of single adds (pend_mu) ------------------------------------------------------------------
    std::vector<uint64_t> pend_keys; std::vector<float> pend_vecs; FlatMap pend_pos;
    // ---- build scratch (mu) --------------------------------------------------------------------------------
    BuildBatch bb{}; uint32_t efpA = 128; DBuf<unsigned char> build_scratch, exact_scratch, tc_scratch, stage; DBuf<uint32_t> tomb, cmap;
    uint32_t* knn_set_id[KNN_SETS][MAX_LEVELS] = {}; float* knn_set_d[KNN_SETS][MAX_LEVELS] = {};
    TcStats tc_build_stats; std::vector<cudaEvent_t> trace_ev;
    // ---- search scratch (search_mu) ------------------------------------------------------------------------
    DBuf<unsigned char> s_exact_scratch, s_tc_scratch; DBuf<float> qpad, qdense, out_d; DBuf<uint64_t> out_k; DBuf<uint32_t> found_d; PinBuf<uint32_t> found_pin;
    TcStats tc_stats; DBuf<uint32_t> tomb1;
    std::vector<uint32_t> tomb_queue;
    // ---- counters ------------------------------------------------------------------------------------------
    unsigned long long* counters = nullptr; std::atomic<bool> stats_on{false};
    std::atomic<uint64_t> st_build_vectors{0}, st_build_batches{0}, st_stream_queries{0}, st_search_calls{0}, st_search_launches{0}, st_reserve_calls{0}, st_compactions{0};
    double st_build_seconds = 0, st_reserve_seconds = 0, st_compact_seconds = 0; size_t device_bytes = 0;

    // ---- lifecycle -----------------------------------------------------------------------------------------
    static vs_status create(const ShardOptions& o, Shard** out);
    void destroy();
    // ---- mutation ------------------------------------------------------------------------------------------
    vs_status add_host(size_t n, const uint64_t* keys, const float* vecs);          // vs_add / vs_add_batch
    vs_status add_device(size_t n, const uint64_t* keys, const float* d_vecs);      // rows in this device's HBM
    vs_status remove(uint64_t key);
    vs_status flush();                                                              // lands every staged add; waits for a running flush
    vs_status reserve(size_t cap);
    size_t count();                                                                 // flush, then live
    // ---- search --------------------------------------------------------------------------------------------
    // device-accessible queries / outputs on stream `st` (never null); sync: wait for completion before returning
    vs_status search_device(size_t nq, const float* d_q, size_t k, size_t ef, int exact, uint64_t* d_keys, float* d_dist, cudaStream_t st, bool sync,
                            const OutFan* fan = nullptr);
    vs_status search_host(size_t nq, const float* qs, size_t k, size_t ef, int exact, uint64_t* keys_out, float* dist_out, size_t* found);
    // ---- misc ----------------------------------------------------------------------------------------------
    vs_status get_stats(vs_stats* o);
    vs_status reset_stats();
    vs_status export_level(uint32_t level, uint32_t* links_out, size_t* n_rows, size_t* row_width, uint64_t* slot_keys_out, size_t* slots);
    vs_status save(const char* path);
    static vs_status load(const char* path, const vs_options* overrides, Shard** out);

    // ---- internals -----------------------------------------------------------------------------------------
    GraphView view(uint32_t visible, int top) const;
    int top_for(uint32_t n) const;
    void snapshot(uint32_t& n, int& top);
    vs_status do_reserve(size_t cap);                          // mu held
    vs_status setup_build_scratch();
    vs_status flush_locked();                                  // mu held
    vs_status land(const uint64_t* keys, const float* rows, size_t n, RowSrc src);   // mu held
    vs_status batch_knn(uint32_t first, uint32_t count, int buf, cudaStream_t ks);
    vs_status batch_search(uint32_t first, uint32_t count, int buf, BuildBatch& b);
    vs_status batch_link(const BuildBatch& b, cudaEvent_t knn_ready);
    void batch_shape(uint32_t first, uint32_t count, BuildBatch& b) const;
    vs_status maybe_compact_locked();                          // mu held
    vs_status compact_locked();                                // mu held
    void maybe_flush_nonblocking();
    vs_status wait_async(cudaStream_t st);
    vs_status apply_tombstones();
    cudaEvent_t take_event();
};

}  // namespace vsann
