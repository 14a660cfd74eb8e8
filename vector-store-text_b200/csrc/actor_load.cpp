// actor_load.cpp — native load generator for the index actor (include/vsactor.h): BASELINE configs[0]
// ("10k x 128-d cosine, add + 1k ANN queries k=10 via the index actor") driven the way the reference's HTTP
// handlers drive it - many client threads, one message per vector / query - without a Python client in the way.
//   actor_load <backend.so> <prefix vs_|ovs_> [n=10000] [d=128] [nq=1000] [k=10] [clients=32] [threads=0] [rounds=5]
// The backend is dlopen-ed and handed to libvsactor.so as a function table. Prints one JSON line.
#include <dlfcn.h>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../../include/vsactor.h"

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: actor_load backend.so prefix [n d nq k clients threads rounds]\n"); return 2; }
    const std::string so = argv[1], pre = argv[2];
    auto arg = [&](int i, long def) { return argc > i ? atol(argv[i]) : def; };
    const size_t n = arg(3, 10000), d = arg(4, 128), nq = arg(5, 1000), k = arg(6, 10), clients = arg(7, 32), threads = arg(8, 0), rounds = arg(9, 5);
    void* h = dlopen(so.c_str(), RTLD_NOW | RTLD_GLOBAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 1; }
    vs_backend be{};
    auto sym = [&](const char* name) { void* p = dlsym(h, (pre + name).c_str()); if (!p) { fprintf(stderr, "missing %s%s\n", pre.c_str(), name); exit(1); } return p; };
    be.index_create = (decltype(be.index_create))sym("index_create"); be.index_destroy = (decltype(be.index_destroy))sym("index_destroy");
    be.add = (decltype(be.add))sym("add"); be.remove = (decltype(be.remove))sym("remove"); be.search = (decltype(be.search))sym("search");
    be.count = (decltype(be.count))sym("count"); be.capacity = (decltype(be.capacity))sym("capacity"); be.reserve = (decltype(be.reserve))sym("reserve");
    be.last_error = (decltype(be.last_error))sym("last_error"); be.add_replaces = 1;

    std::mt19937_64 rng(42); std::normal_distribution<float> nd(0.f, 1.f);
    std::vector<float> x(n * d), q(nq * d);
    for (auto& v : x) v = nd(rng);
    for (auto& v : q) v = nd(rng);

    vs_options o{}; o.dims = d; o.metric = VS_METRIC_COS; o.capacity_hint = n;
    vs_actor* a = nullptr;
    if (vs_actor_spawn(&be, "vector.ann", &o, threads, &a) != VS_OK) { fprintf(stderr, "spawn: %s\n", vs_actor_last_error()); return 1; }

    // ingest: `clients` producers, one AddOrReplace message per vector; done when Count sees them all
    double t0 = now();
    {
        std::atomic<size_t> next{0}; std::vector<std::thread> th;
        for (size_t c = 0; c < clients; ++c) th.emplace_back([&] {
            for (size_t i; (i = next.fetch_add(1)) < n;) { uint64_t pk = i; vs_actor_add_or_replace(a, &pk, sizeof pk, x.data() + i * d, d); }
        });
        for (auto& t : th) t.join();
    }
    size_t cnt = 0;
    while (vs_actor_count(a, &cnt) == VS_OK && cnt != n) std::this_thread::sleep_for(std::chrono::microseconds(200));
    const double t_add = now() - t0;

    auto ann_round = [&](size_t nclients) {
        std::atomic<size_t> next{0}, bad{0}; std::vector<std::thread> th;
        const double s = now();
        for (size_t c = 0; c < nclients; ++c) th.emplace_back([&] {
            for (size_t i; (i = next.fetch_add(1)) < nq;) {
                vs_matches* m = nullptr;
                if (vs_actor_ann(a, q.data() + i * d, d, k, &m) != VS_OK || vs_matches_len(m) != std::min(k, n)) bad.fetch_add(1);
                vs_matches_free(m);
            }
        });
        for (auto& t : th) t.join();
        const double dt = now() - s;
        if (bad.load()) fprintf(stderr, "%zu bad replies\n", bad.load());
        return dt;
    };
    ann_round(1);   // warm-up
    double one = 1e30, many = 1e30;
    for (size_t r = 0; r < rounds; ++r) one = std::min(one, ann_round(1));
    for (size_t r = 0; r < rounds; ++r) many = std::min(many, ann_round(clients));
    vs_actor_stats st{}; vs_actor_get_stats(a, &st);
    printf("{\"backend\": \"%s\", \"n\": %zu, \"d\": %zu, \"nq\": %zu, \"k\": %zu, \"clients\": %zu, \"actor_threads\": %llu, "
           "\"add_vectors_per_s\": %.1f, \"ann_qps_1_client\": %.1f, \"ann_latency_ms_1_client\": %.4f, \"ann_qps_clients\": %.1f, "
           "\"max_in_flight\": %llu, \"max_queued\": %llu}\n",
           so.c_str(), n, d, nq, k, clients, (unsigned long long)st.threads, n / t_add, nq / one, 1e3 * one / nq, nq / many,
           (unsigned long long)st.max_in_flight, (unsigned long long)st.max_queued);
    vs_actor_close(a);
    return 0;
}
