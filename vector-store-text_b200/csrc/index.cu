// index.cu — the C ABI of include/vsann.h over one or several shards (shard.cuh).
//
// One vs_index = one process-local object, as the reference's actor holds one Arc<RwLock<usearch::Index>>
// (src/index/usearch.rs:98,105-136). With vs_options.n_devices in 2..8 it owns one shard per GPU of the box:
//   add / remove     routed by shard = key % G, shards build concurrently, no collective
//   search           the query block is read by every GPU (in place from pinned host memory, or from the serving GPU's HBM over
//                    NVLink), each shard runs K3 / K1 on its own stream, and K6 on the serving GPU (device_ids[0]) merges the G
//                    per-shard lists it reads in place through peer mappings - one launch after the traversals, no all-gather.
//   count            sum of the shard counts
// With one device every call goes straight to the shard.
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "shard.cuh"

using namespace vsann;

struct vs_index {
    int G = 1; Shard* sh[8] = {}; int devs[8] = {};
    uint32_t dims = 0;
    // multi-device search state (msearch_mu)
    std::mutex msearch_mu;
    DBuf<uint64_t> rk[8]; DBuf<float> rd[8];                  // per-shard result lists, on the shard's device, peer-mapped on the serving device
    DBuf<uint64_t> mk; DBuf<float> md; DBuf<float> mq;        // serving device: merged lists, staged queries
    uint64_t** d_pk = nullptr; float** d_pd = nullptr;        // serving device: arrays of G pointers
    uint64_t* h_pk[8] = {}; float* h_pd[8] = {}; size_t ptr_cap = 0;
    cudaEvent_t done[8] = {}, ready = nullptr, merged = nullptr; bool merged_valid = false;   // merged: the last merge has read the shards' result lists
    DBuf<uint32_t> mfound_d; PinBuf<uint32_t> mfound_pin;
    // multi-device add staging: per-shard contiguous copies of the caller's rows
    std::mutex madd_mu; std::vector<uint64_t> part_keys[8]; PinBuf<float> part_rows[8]; size_t part_n[8] = {};
    DBuf<float> dev_part[8];                                  // add_batch_device: rows gathered for shard s, on shard s's device
    DBuf<float> gather_src; DBuf<uint32_t> gather_idx;        // on the serving device
    // vs_search coalescer: single-query calls that arrive while a launch is in flight ride the next launch together
    struct SearchReq { const float* q; size_t k; uint64_t* keys; float* dist; size_t* found; vs_status st; std::string err; bool done; };
    std::mutex co_mu; std::condition_variable co_cv; std::vector<SearchReq*> co_q; bool co_leader = false, coalesce = true;
    PinBuf<float> co_pin_q; PinBuf<uint64_t> co_pin_k; PinBuf<float> co_pin_d;
};

namespace vsann {
__global__ void k_gather_rows_idx(const float* __restrict__ src, float* __restrict__ dst, const uint32_t* __restrict__ idx, uint32_t n, uint32_t dims) {
    const size_t total = (size_t)n * dims;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(i / dims), c = (uint32_t)(i % dims);
        dst[i] = src[(size_t)idx[r] * dims + c];
    }
}
}

namespace {

vs_status msearch_buffers(vs_index* ix, size_t nq, size_t k) {
    const size_t want = nq * k;
    if (want > ix->ptr_cap) {
        for (int s = 0; s < ix->G; ++s) {
            CU(cudaSetDevice(ix->devs[s]));
            CU(ix->rk[s].ensure(want)); CU(ix->rd[s].ensure(want));
            ix->h_pk[s] = ix->rk[s].p; ix->h_pd[s] = ix->rd[s].p;
        }
        CU(cudaSetDevice(ix->devs[0]));
        CU(ix->mk.ensure(want)); CU(ix->md.ensure(want));
        if (!ix->d_pk) { CU(cudaMalloc(&ix->d_pk, 8 * sizeof(void*))); CU(cudaMalloc(&ix->d_pd, 8 * sizeof(void*))); }
        CU(cudaMemcpy(ix->d_pk, ix->h_pk, 8 * sizeof(void*), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(ix->d_pd, ix->h_pd, 8 * sizeof(void*), cudaMemcpyHostToDevice));
        ix->ptr_cap = want;
    }
    return VS_OK;
}

// Fan-out + fused merge. q: a pointer every device can read (pinned host memory, or the serving device's HBM with peer access);
// results land in ix->mk / ix->md on the serving device, on stream `ms` (the serving shard's search stream unless the caller gave one).
vs_status msearch_enqueue(vs_index* ix, size_t nq, const float* q, size_t k, size_t ef, int exact, uint64_t* out_k, float* out_d, cudaStream_t ms, cudaEvent_t after) {
    for (int s = 0; s < ix->G; ++s) {
        Shard* sh = ix->sh[s];
        CU(cudaSetDevice(ix->devs[s]));
        if (after) CU(cudaStreamWaitEvent(sh->sstream, after, 0));
        if (ix->merged_valid) CU(cudaStreamWaitEvent(sh->sstream, ix->merged, 0));    // the previous merge still reads this shard's lists
        vs_status st = sh->search_device(nq, q, k, ef, exact, ix->rk[s].p, ix->rd[s].p, sh->sstream, false); if (st != VS_OK) return st;
        CU(cudaEventRecord(ix->done[s], sh->sstream));
    }
    CU(cudaSetDevice(ix->devs[0]));
    for (int s = 0; s < ix->G; ++s) CU(cudaStreamWaitEvent(ms, ix->done[s], 0));
    CU(launch_merge_topk_ptrs((uint32_t)ix->G, (uint32_t)nq, (uint32_t)k, ix->d_pk, ix->d_pd, out_k, out_d, ms));
    CU(cudaEventRecord(ix->merged, ms)); ix->merged_valid = true;
    return VS_OK;
}

vs_status msearch_host(vs_index* ix, size_t nq, const float* qs, size_t k, size_t ef, int exact, uint64_t* keys_out, float* dist_out, size_t* found) {
    std::lock_guard<std::mutex> g(ix->msearch_mu);
    vs_status s = msearch_buffers(ix, nq, k); if (s != VS_OK) return s;
    CU(cudaSetDevice(ix->devs[0]));
    cudaStream_t ms = ix->sh[0]->sstream;
    // pinned host queries are read in place by every GPU; anything else is staged once on the serving device and read from there over NVLink
    const float* q = nullptr;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, qs) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) q = (const float*)at.devicePointer;
    else cudaGetLastError();
    cudaEvent_t after = nullptr;
    if (!q) {
        CU(ix->mq.ensure(nq * ix->dims));
        CU(cudaMemcpyAsync(ix->mq.p, qs, nq * ix->dims * 4, cudaMemcpyHostToDevice, ms));
        CU(cudaEventRecord(ix->ready, ms)); after = ix->ready;
        q = ix->mq.p;
    }
    s = msearch_enqueue(ix, nq, q, k, ef, exact, ix->mk.p, ix->md.p, ms, after); if (s != VS_OK) return s;
    CU(cudaMemcpyAsync(keys_out, ix->mk.p, nq * k * 8, cudaMemcpyDeviceToHost, ms));
    CU(cudaMemcpyAsync(dist_out, ix->md.p, nq * k * 4, cudaMemcpyDeviceToHost, ms));
    if (found) {
        CU(ix->mfound_d.ensure(nq)); CU(ix->mfound_pin.ensure(std::max<size_t>(nq, 16384)));
        CU(launch_count_found(ix->mk.p, (uint32_t)nq, (uint32_t)k, ix->mfound_d.p, ms));
        CU(cudaMemcpyAsync(ix->mfound_pin.p, ix->mfound_d.p, nq * 4, cudaMemcpyDeviceToHost, ms));
    }
    CU(cudaStreamSynchronize(ms));
    if (found) for (size_t i = 0; i < nq; ++i) found[i] = ix->mfound_pin.p[i];
    return VS_OK;
}

template <class F> vs_status for_each_shard_parallel(vs_index* ix, F f) {
    vs_status st[8]; std::string err[8]; std::vector<std::thread> th;
    for (int s = 0; s < ix->G; ++s) th.emplace_back([&, s] { st[s] = f(s); if (st[s] != VS_OK) err[s] = g_err; });
    for (auto& t : th) t.join();
    for (int s = 0; s < ix->G; ++s) if (st[s] != VS_OK) { g_err = err[s]; return st[s]; }
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
    const int G = o->n_devices > 1 ? o->n_devices : 1;
    if (G > 8) return fail(VS_ERR_INVALID, "create: at most 8 devices");
    auto* ix = new vs_index();
    ix->G = G; ix->dims = (uint32_t)o->dims;
    for (int s = 0; s < G; ++s) {
        ix->devs[s] = G > 1 ? (int)o->device_ids[s] : o->device;
        if (ix->devs[s] < 0 || ix->devs[s] >= ndev) { delete ix; return fail(VS_ERR_INVALID, "create: bad device ordinal"); }
        for (int t = 0; t < s; ++t) if (ix->devs[t] == ix->devs[s]) { delete ix; return fail(VS_ERR_INVALID, "create: a device is listed twice"); }
    }
    if (const char* e = getenv("VSANN_COALESCE")) ix->coalesce = atoi(e) != 0;     // 0: every vs_search call is its own launch
    ShardOptions so;
    so.dims = (uint32_t)o->dims; so.metric = o->metric; so.M = (uint32_t)M;
    so.efA = (uint32_t)(o->expansion_add ? o->expansion_add : 128); so.efS = (uint32_t)(o->expansion_search ? o->expansion_search : 64);
    so.max_batch = (uint32_t)o->max_batch;
    const size_t hint = o->capacity_hint ? o->capacity_hint : 1000000;   // usearch.rs:62
    so.capacity_hint = G > 1 ? std::max<size_t>((hint + G - 1) / G, 1024) : hint;
    vs_status s = VS_OK;
    for (int i = 0; i < G && s == VS_OK; ++i) { so.device = ix->devs[i]; s = Shard::create(so, &ix->sh[i]); }
    if (s == VS_OK && G > 1) {
        auto body = [&]() -> vs_status {
            for (int a = 0; a < G; ++a) {          // every device reads the serving device's query block, the serving device reads every result list
                CU(cudaSetDevice(ix->devs[a]));
                for (int b = 0; b < G; ++b) {
                    if (a == b) continue;
                    int can = 0; CU(cudaDeviceCanAccessPeer(&can, ix->devs[a], ix->devs[b]));
                    if (!can) return fail(VS_ERR_CUDA, "create: devices " + std::to_string(ix->devs[a]) + " and " + std::to_string(ix->devs[b]) + " are not peers");
                    const cudaError_t e = cudaDeviceEnablePeerAccess(ix->devs[b], 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU(e);
                    cudaGetLastError();
                }
                CU(cudaEventCreateWithFlags(&ix->done[a], cudaEventDisableTiming));
            }
            CU(cudaSetDevice(ix->devs[0]));
            CU(cudaEventCreateWithFlags(&ix->ready, cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&ix->merged, cudaEventDisableTiming));
            return VS_OK;
        };
        s = body();
    }
    if (s != VS_OK) { std::string keep = g_err; vs_index_destroy(ix); g_err = keep; return s; }
    *out = ix; return VS_OK;
}

void vs_index_destroy(vs_index* ix) {
    if (!ix) return;
    for (int s = 0; s < ix->G; ++s) {
        if (!ix->sh[s]) continue;
        cudaSetDevice(ix->devs[s]);
        cudaStreamSynchronize(ix->sh[s]->sstream);
        ix->rk[s].release(); ix->rd[s].release(); ix->dev_part[s].release(); ix->part_rows[s].release();
        if (ix->done[s]) cudaEventDestroy(ix->done[s]);
    }
    cudaSetDevice(ix->devs[0]);
    ix->mk.release(); ix->md.release(); ix->mq.release(); ix->mfound_d.release(); ix->mfound_pin.release(); ix->gather_src.release(); ix->gather_idx.release();
    if (ix->d_pk) cudaFree(ix->d_pk); if (ix->d_pd) cudaFree(ix->d_pd);
    if (ix->ready) cudaEventDestroy(ix->ready); if (ix->merged) cudaEventDestroy(ix->merged);
    ix->co_pin_q.release(); ix->co_pin_k.release(); ix->co_pin_d.release();
    for (int s = 0; s < ix->G; ++s) if (ix->sh[s]) ix->sh[s]->destroy();
    cudaGetLastError();
    delete ix;
}

vs_status vs_add_batch(vs_index* ix, size_t n, const uint64_t* keys, const float* vecs) {
    if (!ix || (n && (!keys || !vecs))) return fail(VS_ERR_INVALID, "add: null argument");
    if (ix->G == 1) return ix->sh[0]->add_host(n, keys, vecs);
    std::lock_guard<std::mutex> g(ix->madd_mu);
    const size_t d = ix->dims; const int G = ix->G;
    for (int s = 0; s < G; ++s) { ix->part_keys[s].clear(); ix->part_n[s] = 0; }
    for (size_t i = 0; i < n; ++i) ix->part_n[keys[i] % G]++;
    for (int s = 0; s < G; ++s) { CU(ix->part_rows[s].ensure(std::max<size_t>(ix->part_n[s], 1) * d)); ix->part_keys[s].reserve(ix->part_n[s]); }
    for (size_t i = 0; i < n; ++i) {
        const int s = (int)(keys[i] % G);
        std::memcpy(ix->part_rows[s].p + ix->part_keys[s].size() * d, vecs + i * d, d * sizeof(float));
        ix->part_keys[s].push_back(keys[i]);
    }
    if (n < 64) {   // a trickle: no point in spawning threads
        for (int s = 0; s < G; ++s) { vs_status st = ix->sh[s]->add_host(ix->part_keys[s].size(), ix->part_keys[s].data(), ix->part_rows[s].p); if (st != VS_OK) return st; }
        return VS_OK;
    }
    return for_each_shard_parallel(ix, [&](int s) { return ix->sh[s]->add_host(ix->part_keys[s].size(), ix->part_keys[s].data(), ix->part_rows[s].p); });
}
vs_status vs_add(vs_index* ix, uint64_t key, const float* vec) {
    if (!ix || !vec) return fail(VS_ERR_INVALID, "add: null argument");
    return ix->sh[key % ix->G]->add_host(1, &key, vec);
}

vs_status vs_add_batch_device(vs_index* ix, size_t n, const uint64_t* keys, const float* d_vecs) {
    if (!ix || (n && (!keys || !d_vecs))) return fail(VS_ERR_INVALID, "add: null argument");
    if (ix->G == 1) return ix->sh[0]->add_device(n, keys, d_vecs);
    // rows live on the serving device: gather each shard's rows there, move them over NVLink, land them shard by shard in parallel
    std::lock_guard<std::mutex> g(ix->madd_mu);
    const size_t d = ix->dims; const int G = ix->G;
    std::vector<uint32_t> idx[8];
    for (int s = 0; s < G; ++s) ix->part_keys[s].clear();
    for (size_t i = 0; i < n; ++i) { const int s = (int)(keys[i] % G); idx[s].push_back((uint32_t)i); ix->part_keys[s].push_back(keys[i]); }
    CU(cudaSetDevice(ix->devs[0]));
    CU(cudaStreamSynchronize(nullptr));
    cudaStream_t ms = ix->sh[0]->sstream;
    size_t biggest = 0; for (int s = 0; s < G; ++s) biggest = std::max(biggest, idx[s].size());
    CU(ix->gather_src.ensure(std::max<size_t>(biggest, 1) * d)); CU(ix->gather_idx.ensure(std::max<size_t>(biggest, 1)));
    for (int s = 0; s < G; ++s) {
        const size_t m = idx[s].size();
        if (!m) continue;
        CU(cudaSetDevice(ix->devs[s])); CU(ix->dev_part[s].ensure(m * d));
        CU(cudaSetDevice(ix->devs[0]));
        CU(cudaMemcpyAsync(ix->gather_idx.p, idx[s].data(), m * 4, cudaMemcpyHostToDevice, ms));
        const uint32_t blocks = (uint32_t)std::min<size_t>((m * d + 255) / 256, 148 * 16);
        vsann::k_gather_rows_idx<<<blocks, 256, 0, ms>>>(d_vecs, ix->gather_src.p, ix->gather_idx.p, (uint32_t)m, (uint32_t)d);
        CU(cudaGetLastError());
        CU(cudaMemcpyPeerAsync(ix->dev_part[s].p, ix->devs[s], ix->gather_src.p, ix->devs[0], m * d * 4, ms));
        CU(cudaStreamSynchronize(ms));
    }
    return for_each_shard_parallel(ix, [&](int s) -> vs_status {
        if (ix->part_keys[s].empty()) return VS_OK;
        cudaSetDevice(ix->devs[s]);
        return ix->sh[s]->add_device(ix->part_keys[s].size(), ix->part_keys[s].data(), ix->dev_part[s].p);
    });
}

vs_status vs_remove(vs_index* ix, uint64_t key) {
    if (!ix) return fail(VS_ERR_INVALID, "remove: null index");
    return ix->sh[key % ix->G]->remove(key);
}

vs_status vs_search_batch_device(vs_index* ix, size_t nq, const float* d_q, size_t k, size_t ef, int exact,
                                 uint64_t* d_keys, float* d_dist, void* stream) {
    if (!ix || (nq && (!d_q || !d_keys || !d_dist))) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    if (nq == 0) return VS_OK;
    if (ix->G == 1) {
        Shard* sh = ix->sh[0];
        return sh->search_device(nq, d_q, k, ef, exact, d_keys, d_dist, stream ? (cudaStream_t)stream : sh->sstream, stream == nullptr);
    }
    // queries / outputs on the serving device (or pinned host memory); every shard reads the queries over NVLink
    std::lock_guard<std::mutex> g(ix->msearch_mu);
    vs_status s = msearch_buffers(ix, nq, k); if (s != VS_OK) return s;
    CU(cudaSetDevice(ix->devs[0]));
    cudaStream_t ms = stream ? (cudaStream_t)stream : ix->sh[0]->sstream;
    CU(cudaEventRecord(ix->ready, ms));
    s = msearch_enqueue(ix, nq, d_q, k, ef, exact, d_keys, d_dist, ms, ix->ready); if (s != VS_OK) return s;
    if (!stream) CU(cudaStreamSynchronize(ms));
    return VS_OK;
}

vs_status vs_search_batch_fanout_device(vs_index* ix, size_t nq, const float* d_q, size_t k, size_t ef, size_t n_out,
                                        uint64_t* const* d_keys, float* const* d_dist, void* stream) {
    if (!ix || (nq && (!d_q || !d_keys || !d_dist))) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    if (n_out == 0 || n_out > 9) return fail(VS_ERR_UNSUPPORTED, "search: 1 to 9 result buffers");
    if (ix->G != 1) return fail(VS_ERR_UNSUPPORTED, "search: result fan-out is for single-device indexes (a multi-device index merges inside the library)");
    for (size_t i = 0; i < n_out; ++i) if (nq && (!d_keys[i] || !d_dist[i])) return fail(VS_ERR_INVALID, "search: null result buffer");
    if (nq == 0) return VS_OK;
    OutFan fan; fan.n = (uint32_t)(n_out - 1);
    for (size_t i = 1; i < n_out; ++i) { fan.keys[i - 1] = d_keys[i]; fan.dist[i - 1] = d_dist[i]; }
    Shard* sh = ix->sh[0];
    return sh->search_device(nq, d_q, k, ef, 0, d_keys[0], d_dist[0], stream ? (cudaStream_t)stream : sh->sstream, stream == nullptr, &fan);
}

vs_status vs_search_batch(vs_index* ix, size_t nq, const float* qs, size_t k, size_t ef, int exact,
                          uint64_t* keys_out, float* dist_out, size_t* found) {
    if (!ix || (nq && (!qs || !keys_out || !dist_out))) return fail(VS_ERR_INVALID, "search: null argument");
    if (k == 0) return fail(VS_ERR_INVALID, "search: k == 0");
    if (nq == 0) return VS_OK;
    if (ix->G == 1) return ix->sh[0]->search_host(nq, qs, k, ef, exact, keys_out, dist_out, found);
    return msearch_host(ix, nq, qs, k, ef, exact, keys_out, dist_out, found);
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
        if (ix->co_pin_q.ensure(std::max<size_t>(n, 256) * d) != cudaSuccess || ix->co_pin_k.ensure(std::max<size_t>(n, 256) * k) != cudaSuccess ||
            ix->co_pin_d.ensure(std::max<size_t>(n, 256) * k) != cudaSuccess) { cudaGetLastError(); s = fail(VS_ERR_NOMEM, "search: pinned staging allocation failed"); }
        if (s == VS_OK) {
            for (size_t i = 0; i < n; ++i) std::memcpy(ix->co_pin_q.p + i * d, grp[i]->q, d * sizeof(float));
            s = vs_search_batch(ix, n, ix->co_pin_q.p, k, 0, 0, ix->co_pin_k.p, ix->co_pin_d.p, nullptr);
        }
        for (size_t i = 0; i < n; ++i) {
            vs_index::SearchReq* r = grp[i];
            r->st = s;
            if (s != VS_OK) { r->err = g_err; continue; }
            std::memcpy(r->keys, ix->co_pin_k.p + i * k, k * 8); std::memcpy(r->dist, ix->co_pin_d.p + i * k, k * 4);
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
    size_t c = 0;
    for (int s = 0; s < ix->G; ++s) c += ix->sh[s]->count();
    return c;
}
size_t vs_capacity(vs_index* ix) { if (!ix) return 0; size_t c = 0; for (int s = 0; s < ix->G; ++s) c += ix->sh[s]->cap_pub.load(); return c; }
vs_status vs_reserve(vs_index* ix, size_t cap) {
    if (!ix) return fail(VS_ERR_INVALID, "reserve: null index");
    for (int s = 0; s < ix->G; ++s) { vs_status st = ix->sh[s]->reserve((cap + ix->G - 1) / ix->G); if (st != VS_OK) return st; }
    return VS_OK;
}
vs_status vs_flush(vs_index* ix) {
    if (!ix) return fail(VS_ERR_INVALID, "flush: null index");
    if (ix->G == 1) return ix->sh[0]->flush();
    return for_each_shard_parallel(ix, [&](int s) { return ix->sh[s]->flush(); });
}

vs_status vs_get_stats(vs_index* ix, vs_stats* o) {
    if (!ix || !o) return fail(VS_ERR_INVALID, "stats: null argument");
    vs_status s = ix->sh[0]->get_stats(o); if (s != VS_OK) return s;
    o->n_devices = (uint32_t)ix->G;
    for (int i = 1; i < ix->G; ++i) {
        vs_stats t; s = ix->sh[i]->get_stats(&t); if (s != VS_OK) return s;
        o->count += t.count; o->slots += t.slots; o->capacity += t.capacity; o->pending += t.pending;
        o->search_queries += t.search_queries; o->search_dist_evals += t.search_dist_evals; o->search_expansions += t.search_expansions;
        o->build_vectors += t.build_vectors; o->build_dist_evals += t.build_dist_evals; o->build_batches += t.build_batches; o->build_expansions += t.build_expansions;
        o->build_seconds = std::max(o->build_seconds, t.build_seconds); o->device_bytes += t.device_bytes; o->top_level = std::max(o->top_level, t.top_level);
        o->exact_tc_queries += t.exact_tc_queries; o->exact_tc_fallback += t.exact_tc_fallback; o->exact_tc_second_scan += t.exact_tc_second_scan;
        o->exact_tc_scans += t.exact_tc_scans; o->exact_tc_scan_seconds = std::max(o->exact_tc_scan_seconds, t.exact_tc_scan_seconds);
        o->exact_stream_queries += t.exact_stream_queries; o->reserve_calls += t.reserve_calls; o->reserve_seconds = std::max(o->reserve_seconds, t.reserve_seconds);
        o->compactions += t.compactions; o->compact_seconds = std::max(o->compact_seconds, t.compact_seconds);
    }
    return VS_OK;
}
vs_status vs_set_stats(vs_index* ix, int on) {
    if (!ix) return VS_ERR_INVALID;
    for (int s = 0; s < ix->G; ++s) { ix->sh[s]->stats_on.store(on != 0); std::lock_guard<std::mutex> g(ix->sh[s]->search_mu); ix->sh[s]->tc_stats.time_scan = on != 0; }
    return VS_OK;
}
vs_status vs_reset_stats(vs_index* ix) {
    if (!ix) return VS_ERR_INVALID;
    for (int s = 0; s < ix->G; ++s) { vs_status st = ix->sh[s]->reset_stats(); if (st != VS_OK) return st; }
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

vs_status vs_merge_topk_flags_device(int device, size_t parts, size_t nq, size_t k, const uint64_t* const* d_part_keys, const float* const* d_part_dist,
                                     uint32_t* const* d_part_flags, uint32_t my_part, uint32_t epoch,
                                     uint64_t* d_keys_out, float* d_dist_out, void* stream) {
    if (!d_part_keys || !d_part_dist || !d_part_flags || !d_keys_out || !d_dist_out) return fail(VS_ERR_INVALID, "merge: null argument");
    CU(cudaSetDevice(device));
    CU(launch_merge_topk_flags((uint32_t)parts, (uint32_t)nq, (uint32_t)k, d_part_keys, d_part_dist, d_part_flags, my_part, epoch, d_keys_out, d_dist_out, (cudaStream_t)stream));
    if (!stream) CU(cudaStreamSynchronize(nullptr));
    return VS_OK;
}

vs_status vs_export_level(vs_index* ix, uint32_t level, uint32_t* links_out, size_t* n_rows, size_t* row_width, uint64_t* slot_keys_out, size_t* slots) {
    if (!ix || !n_rows || !row_width) return fail(VS_ERR_INVALID, "export: null argument");
    if (ix->G > 1) return fail(VS_ERR_UNSUPPORTED, "export: one shard at a time (single-device indexes only)");
    return ix->sh[0]->export_level(level, links_out, n_rows, row_width, slot_keys_out, slots);
}

// Multi-device indexes are saved as one file per shard: <path>.<shard>of<G>.
vs_status vs_index_save(vs_index* ix, const char* path) {
    if (!ix || !path) return fail(VS_ERR_INVALID, "save: null argument");
    if (ix->G == 1) return ix->sh[0]->save(path);
    for (int s = 0; s < ix->G; ++s) {
        const std::string p = std::string(path) + "." + std::to_string(s) + "of" + std::to_string(ix->G);
        vs_status st = ix->sh[s]->save(p.c_str()); if (st != VS_OK) return st;
    }
    return VS_OK;
}

vs_status vs_index_load(const char* path, const vs_options* ov, vs_index** out) {
    if (!path || !out) return fail(VS_ERR_INVALID, "load: null argument");
    *out = nullptr;
    const int G = ov && ov->n_devices > 1 ? ov->n_devices : 1;
    if (G > 1) return fail(VS_ERR_UNSUPPORTED, "load: multi-device indexes are restored shard by shard (load <path>.<s>of<G> into single-device indexes)");
    Shard* sh = nullptr;
    vs_status s = Shard::load(path, ov, &sh); if (s != VS_OK) return s;
    auto* ix = new vs_index();
    ix->G = 1; ix->sh[0] = sh; ix->devs[0] = sh->dev.device; ix->dims = sh->dims;
    if (const char* e = getenv("VSANN_COALESCE")) ix->coalesce = atoi(e) != 0;
    *out = ix; return VS_OK;
}

const char* vs_last_error(void) { return g_err.c_str(); }
const char* vs_version(void) { return "vsann-b200 0.2 sm_100a"; }

}  // extern "C"
