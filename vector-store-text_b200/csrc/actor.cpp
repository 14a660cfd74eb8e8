// actor.cpp — the index actor of src/index/usearch.rs:82-311 as a C++ shim over include/vsann.h
// (include/vsactor.h). Host code only: the compute is whatever backend the
// caller hands over (libvsann.so on the GPU, the CPU oracle in the CPU tests).
//
//   channel(10)                      usearch.rs:102-103   -> Channel (bounded, blocking send)
//   semaphore 2 x threads            usearch.rs:118       -> in_flight permits
//   rayon pool of T threads          usearch.rs:200,244,275 -> worker threads
//   BiMap<PrimaryKey, Key> + AtomicU64 key allocator  usearch.rs:110-113,181-196
//   reserve threshold / increment    usearch.rs:62-66,201-212
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/vsactor.h"

namespace {

constexpr size_t CHANNEL_SIZE = 10;             // usearch.rs:102
constexpr size_t RESERVE_INCREMENT = 1000000;   // usearch.rs:62
constexpr size_t RESERVE_THRESHOLD = RESERVE_INCREMENT / 3;   // usearch.rs:66

thread_local std::string g_actor_err;

struct Reply {                 // oneshot channel
    std::mutex mu; std::condition_variable cv; bool done = false;
    vs_status status = VS_OK; std::string error;
    std::vector<std::string> pks; std::vector<float> distances; size_t count = 0;
    void send() { { std::lock_guard<std::mutex> g(mu); done = true; } cv.notify_all(); }
    void wait() { std::unique_lock<std::mutex> l(mu); cv.wait(l, [&] { return done; }); }
};

struct Message {               // enum Index — usearch.rs:148-170
    enum Kind { AddOrReplace, Remove, Ann, Count } kind;
    std::string pk; std::vector<float> embedding; size_t limit = 0;
    std::shared_ptr<Reply> tx;
};

}  // namespace

struct vs_matches { std::vector<std::string> pks; std::vector<float> distances; };

struct vs_actor {
    vs_backend be{}; vs_index* idx = nullptr; std::string id; size_t dims = 0, threads = 1;
    // channel
    std::mutex ch_mu; std::condition_variable ch_not_full, ch_not_empty; std::deque<Message> channel; bool closed = false;
    // in-flight permits (semaphore 2 x T) + the pool's task queue
    std::mutex pool_mu; std::condition_variable pool_cv, permit_cv; std::deque<Message> tasks; size_t in_flight = 0; bool pool_stop = false;
    std::thread consumer; std::vector<std::thread> workers;
    // bimap + key allocator
    std::mutex keys_mu; std::unordered_map<std::string, uint64_t> by_pk; std::unordered_map<uint64_t, std::string> by_key;
    std::atomic<uint64_t> next_key{0};
    // stats
    std::atomic<uint64_t> n_messages{0}, n_rollbacks{0}; uint64_t max_in_flight = 0, max_queued = 0;

    void consume();
    void work();
    void process(Message& m);
    void add_or_replace(Message& m);
    void remove(Message& m);
    void ann(Message& m);
};

// `while let Some(msg) = rx.recv().await { permit = semaphore.acquire_owned().await; spawn(process) }` — usearch.rs:120-131
void vs_actor::consume() {
    for (;;) {
        Message m;
        {
            std::unique_lock<std::mutex> l(ch_mu);
            ch_not_empty.wait(l, [&] { return !channel.empty() || closed; });
            if (channel.empty()) break;                   // closed and drained
            m = std::move(channel.front()); channel.pop_front();
        }
        ch_not_full.notify_one();
        {
            std::unique_lock<std::mutex> l(pool_mu);
            permit_cv.wait(l, [&] { return in_flight < 2 * threads; });
            ++in_flight; if (in_flight > max_in_flight) max_in_flight = in_flight;
            tasks.push_back(std::move(m));
        }
        pool_cv.notify_one();
    }
    { std::lock_guard<std::mutex> l(pool_mu); pool_stop = true; }
    pool_cv.notify_all();
}

void vs_actor::work() {
    for (;;) {
        Message m;
        {
            std::unique_lock<std::mutex> l(pool_mu);
            pool_cv.wait(l, [&] { return !tasks.empty() || pool_stop; });
            if (tasks.empty()) return;
            m = std::move(tasks.front()); tasks.pop_front();
        }
        process(m);
        n_messages.fetch_add(1);
        { std::lock_guard<std::mutex> l(pool_mu); --in_flight; }   // drop(permit)
        permit_cv.notify_one();
    }
}

void vs_actor::process(Message& m) {   // usearch.rs:141-172
    switch (m.kind) {
        case Message::AddOrReplace: add_or_replace(m); break;
        case Message::Remove: remove(m); break;
        case Message::Ann: ann(m); break;
        case Message::Count: m.tx->count = be.count(idx); m.tx->send(); break;   // usearch.rs:308-311
    }
}

// usearch.rs:174-233
void vs_actor::add_or_replace(Message& m) {
    uint64_t key = next_key.fetch_add(1);
    bool replace = false;
    {
        std::lock_guard<std::mutex> g(keys_mu);
        auto it = by_pk.find(m.pk);
        if (it == by_pk.end()) { by_pk.emplace(m.pk, key); by_key.emplace(key, m.pk); }   // insert_no_overwrite
        else { next_key.fetch_sub(1); key = it->second; replace = true; }
    }
    bool ok = true;
    // capacity - size < RESERVE_THRESHOLD => reserve(capacity + RESERVE_INCREMENT) — usearch.rs:201-212.
    // `size` is the bimap's: asking the library would force staged adds to land one by one.
    size_t size;
    { std::lock_guard<std::mutex> g(keys_mu); size = by_pk.size(); }
    const size_t capacity = be.capacity(idx);
    if (capacity < size || capacity - size < RESERVE_THRESHOLD) {
        if (be.reserve(idx, capacity + RESERVE_INCREMENT) != VS_OK) ok = false;
    }
    if (ok && replace && !be.add_replaces) {
        const vs_status s = be.remove(idx, key);
        if (s != VS_OK && s != VS_NOT_FOUND) return;          // "don't remove a key from a bimap" — usearch.rs:216-219
    }
    if (ok && be.add(idx, key, m.embedding.data()) != VS_OK) ok = false;
    if (!ok) {                                                  // usearch.rs:230-232
        std::lock_guard<std::mutex> g(keys_mu);
        auto it = by_key.find(key);
        if (it != by_key.end()) { by_pk.erase(it->second); by_key.erase(it); }
        n_rollbacks.fetch_add(1);
    }
}

// usearch.rs:235-249
void vs_actor::remove(Message& m) {
    uint64_t key;
    {
        std::lock_guard<std::mutex> g(keys_mu);
        auto it = by_pk.find(m.pk);
        if (it == by_pk.end()) return;
        key = it->second; by_key.erase(key); by_pk.erase(it);
    }
    be.remove(idx, key);   // failures are only logged by the reference
}

// usearch.rs:251-306
void vs_actor::ann(Message& m) {
    Reply& r = *m.tx;
    if (m.embedding.empty()) { r.status = VS_ERR_DIMS; r.error = "ann: embedding dimensions == 0"; r.send(); return; }
    if (m.embedding.size() != dims) {
        r.status = VS_ERR_DIMS;
        r.error = "ann: wrong embedding dimensions: " + std::to_string(m.embedding.size()) + " != " + std::to_string(dims);
        r.send(); return;
    }
    std::vector<uint64_t> keys(m.limit); std::vector<float> dist(m.limit); size_t found = 0;
    const vs_status s = be.search(idx, m.embedding.data(), m.limit, keys.data(), dist.data(), &found);
    if (s != VS_OK) { r.status = s; r.error = std::string("ann: search failed: ") + (be.last_error ? be.last_error() : ""); r.send(); return; }
    {
        std::lock_guard<std::mutex> g(keys_mu);
        for (size_t i = 0; i < found; ++i) {
            auto it = by_key.find(keys[i]);
            if (it == by_key.end()) {
                r.status = VS_NOT_FOUND; r.error = "not defined primary key column " + std::to_string(keys[i]);
                r.pks.clear(); r.distances.clear(); r.send(); return;
            }
            r.pks.push_back(it->second); r.distances.push_back(dist[i]);
        }
    }
    r.send();
}

namespace {
vs_status fail(vs_status s, const std::string& m) { g_actor_err = m; return s; }

// tx.send(msg).await: waits while the channel holds CHANNEL_SIZE messages
vs_status send(vs_actor* a, Message&& m) {
    {
        std::unique_lock<std::mutex> l(a->ch_mu);
        if (a->closed) return fail(VS_ERR_INVALID, "actor is closed");
        a->ch_not_full.wait(l, [&] { return a->channel.size() < CHANNEL_SIZE; });
        a->channel.push_back(std::move(m));
        if (a->channel.size() > a->max_queued) a->max_queued = a->channel.size();
    }
    a->ch_not_empty.notify_one();
    return VS_OK;
}
}  // namespace

extern "C" {

vs_status vs_actor_spawn(const vs_backend* be, const char* index_id, const vs_options* o, size_t threads, vs_actor** out) {
    if (!be || !o || !out || !be->index_create || !be->index_destroy || !be->add || !be->remove || !be->search || !be->count || !be->capacity || !be->reserve)
        return fail(VS_ERR_INVALID, "spawn: null argument or incomplete backend");
    *out = nullptr;
    if (threads == 0) {
        if (const char* e = getenv("SCYLLA_USEARCH_BACKGROUND_THREADS")) threads = (size_t)atol(e);   // README.md:14-15
        if (threads == 0) threads = std::thread::hardware_concurrency();
        if (threads == 0) threads = 1;
    }
    std::unique_ptr<vs_actor> a(new vs_actor());
    a->be = *be; a->id = index_id ? index_id : ""; a->dims = o->dims; a->threads = threads;
    vs_options opt = *o;
    if (opt.capacity_hint == 0) opt.capacity_hint = RESERVE_INCREMENT;     // reserve(RESERVE_INCREMENT) — usearch.rs:99
    const vs_status s = be->index_create(&opt, &a->idx);
    if (s != VS_OK) return fail(s, std::string("spawn: ") + (be->last_error ? be->last_error() : "index_create failed"));
    vs_actor* p = a.release();
    for (size_t i = 0; i < threads; ++i) p->workers.emplace_back([p] { p->work(); });
    p->consumer = std::thread([p] { p->consume(); });
    *out = p;
    return VS_OK;
}

void vs_actor_close(vs_actor* a) {
    if (!a) return;
    { std::lock_guard<std::mutex> l(a->ch_mu); a->closed = true; }
    a->ch_not_empty.notify_all(); a->ch_not_full.notify_all();
    if (a->consumer.joinable()) a->consumer.join();
    for (auto& w : a->workers) if (w.joinable()) w.join();
    if (a->idx) a->be.index_destroy(a->idx);
    delete a;
}

vs_status vs_actor_add_or_replace(vs_actor* a, const void* pk, size_t pk_len, const float* emb, size_t dims) {
    if (!a || (!pk && pk_len) || (!emb && dims)) return fail(VS_ERR_INVALID, "add_or_replace: null argument");
    // the reference's Embedding carries its own length and usearch rejects a wrong one; the C ABI takes a
    // bare pointer, so the check happens here and the message is dropped like a failed add
    if (dims != a->dims) return fail(VS_ERR_DIMS, "add_or_replace: wrong embedding dimensions: " + std::to_string(dims) + " != " + std::to_string(a->dims));
    Message m; m.kind = Message::AddOrReplace; m.pk.assign((const char*)pk, pk_len); m.embedding.assign(emb, emb + dims);
    return send(a, std::move(m));
}

vs_status vs_actor_remove(vs_actor* a, const void* pk, size_t pk_len) {
    if (!a || (!pk && pk_len)) return fail(VS_ERR_INVALID, "remove: null argument");
    Message m; m.kind = Message::Remove; m.pk.assign((const char*)pk, pk_len);
    return send(a, std::move(m));
}

vs_status vs_actor_ann(vs_actor* a, const float* emb, size_t dims, size_t limit, vs_matches** out) {
    if (!a || !out || (!emb && dims)) return fail(VS_ERR_INVALID, "ann: null argument");
    *out = nullptr;
    if (limit == 0) return fail(VS_ERR_INVALID, "ann: limit must be non-zero");      // Limit(NonZeroUsize) — lib.rs:235
    Message m; m.kind = Message::Ann; m.embedding.assign(emb, emb + dims); m.limit = limit; m.tx = std::make_shared<Reply>();
    std::shared_ptr<Reply> rx = m.tx;
    const vs_status s = send(a, std::move(m));
    if (s != VS_OK) return s;
    rx->wait();
    if (rx->status != VS_OK) return fail(rx->status, rx->error);
    vs_matches* r = new vs_matches(); r->pks = std::move(rx->pks); r->distances = std::move(rx->distances);
    *out = r;
    return VS_OK;
}

vs_status vs_actor_count(vs_actor* a, size_t* out) {
    if (!a || !out) return fail(VS_ERR_INVALID, "count: null argument");
    Message m; m.kind = Message::Count; m.tx = std::make_shared<Reply>();
    std::shared_ptr<Reply> rx = m.tx;
    const vs_status s = send(a, std::move(m));
    if (s != VS_OK) return s;
    rx->wait();
    *out = rx->count;
    return VS_OK;
}

size_t vs_matches_len(const vs_matches* m) { return m ? m->pks.size() : 0; }
const void* vs_matches_pk(const vs_matches* m, size_t i, size_t* len) {
    if (!m || i >= m->pks.size()) { if (len) *len = 0; return nullptr; }
    if (len) *len = m->pks[i].size();
    return m->pks[i].data();
}
float vs_matches_distance(const vs_matches* m, size_t i) { return (m && i < m->distances.size()) ? m->distances[i] : 0.f; }
void vs_matches_free(vs_matches* m) { delete m; }

vs_status vs_actor_get_stats(vs_actor* a, vs_actor_stats* o) {
    if (!a || !o) return fail(VS_ERR_INVALID, "stats: null argument");
    std::memset(o, 0, sizeof(*o));
    o->messages = a->n_messages.load(); o->rollbacks = a->n_rollbacks.load(); o->next_key = a->next_key.load(); o->threads = a->threads;
    { std::lock_guard<std::mutex> l(a->pool_mu); o->max_in_flight = a->max_in_flight; }
    { std::lock_guard<std::mutex> l(a->ch_mu); o->max_queued = a->max_queued; }
    { std::lock_guard<std::mutex> l(a->keys_mu); o->keys = a->by_pk.size(); }
    return VS_OK;
}

const char* vs_actor_last_error(void) { return g_actor_err.c_str(); }

}  // extern "C"
