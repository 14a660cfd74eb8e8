"""Ingest benchmark of the path in front of the index: an out-of-order DbEmbedding stream (30 % replaces,
5 % deletes, shuffled delivery) -> ItemsMonitor (last-write-wins, src/monitor_items.rs:56-80) -> the native index actor
(libvsactor.so: channel 10, 2 x T in flight, src/index/usearch.rs:102,118-131) -> backend. Reports accepted messages/s,
vectors/s landed, and visible-after latency (time from the last message to count() == live). GPU backend, and the CPU
restatement beside it on a smaller stream (it ingests ~100x slower).
  python scripts/ingest_bench.py [--n 200000] [--d 128] [--cpu-n 20000] [--threads 0]"""
import argparse, json, os, sys, threading, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vector_store_text_b200 as vst
from vector_store_text_b200.service import DbEmbedding, ItemsMonitor


def stream(nkeys, d, seed):
    rng = np.random.default_rng(seed)
    nmsg = int(nkeys * 1.35)
    rows = rng.standard_normal((nmsg, d), dtype=np.float32)
    pk = np.concatenate([np.arange(nkeys), rng.integers(0, nkeys, nmsg - nkeys)])
    ts = rng.permutation(nmsg) + 1
    is_del = rng.random(nmsg) < 0.05
    order = rng.permutation(nmsg)
    newest = {}
    for i in range(nmsg):
        k = int(pk[i])
        if k not in newest or ts[newest[k]] < ts[i]:
            newest[k] = i
    live = sum(1 for i in newest.values() if not is_del[i])
    return rows, pk, ts, is_del, order, live


def run(lib, label, nkeys, d, threads, producers=4):
    rows, pk, ts, is_del, order, live = stream(nkeys, d, 11)
    actor = vst.NativeActor("ks.items", d, metric=vst.Metric.L2SQ, capacity_hint=max(nkeys, 1000), threads=threads, lib=lib)
    mon = ItemsMonitor(actor)
    lock = threading.Lock()          # the reference's monitor is one task: the LWW map is not shared

    def feed(part):
        for i in part:
            with lock:
                mon.add(DbEmbedding((int(pk[i]),), None if is_del[i] else rows[i], int(ts[i])))
    t0 = time.perf_counter()
    parts = np.array_split(order, producers)
    th = [threading.Thread(target=feed, args=(p,)) for p in parts]
    for t in th: t.start()
    for t in th: t.join()
    t_sent = time.perf_counter()
    # done = every message processed (actor counter) and the count has settled. With T > 1 workers the actor - like the reference -
    # does not order messages in flight, so the final count may differ from the LWW model's by a few keys: reported, not required.
    sent = len(order) - mon.dropped
    last, stable_since = -1, time.perf_counter()
    while True:
        c = actor.count()
        now = time.perf_counter()
        if c != last: last, stable_since = c, now
        if c == live or (actor.stats()["messages"] >= sent and now - stable_since > 0.05) or now - t_sent > 120: break
        time.sleep(0.002)
    t_vis = time.perf_counter()
    out = {"backend": label, "keys": nkeys, "messages": len(order), "dims": d, "accepted": len(order) - mon.dropped, "dropped_by_lww": mon.dropped,
           "live_at_end": live, "count_at_end": actor.count(), "messages_per_s": len(order) / (t_sent - t0),
           "vectors_landed_per_s": (len(order) - mon.dropped) / (t_vis - t0), "visible_after_last_message_ms": 1e3 * (t_vis - t_sent),
           "actor_threads": threads or (os.cpu_count() or 1)}
    q = rows[order[:100]]
    t1 = time.perf_counter()
    for v in q: actor.ann(v, 10)
    out["ann_latency_ms_after_ingest"] = 1e3 * (time.perf_counter() - t1) / len(q)
    actor.close()
    return out


if __name__ == "__main__":
    p = argparse.ArgumentParser()
    p.add_argument("--n", type=int, default=200000); p.add_argument("--d", type=int, default=128)
    p.add_argument("--cpu-n", type=int, default=20000); p.add_argument("--threads", type=int, default=0)
    a = p.parse_args()
    res = {"gpu": run(vst.load_library(), "libvsann.so (B200)", a.n, a.d, a.threads)}
    os.environ["OVS_SIMD"] = "1"
    olib = vst.load_library(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "liboracle_vsann.so"), prefix="ovs_")
    res["cpu"] = run(olib, "oracle (CPU restatement, AVX2 distances)", a.cpu_n, a.d, a.threads)
    print(json.dumps(res))
