"""BASELINE configs[0] through the index actor: 10k x 128-d cosine, add + 1k ANN queries k=10, one message per vector /
query through libvsactor.so (csrc/actor.cpp: channel 10, 2 x T in flight, PK<->key bimap; SURVEY.md §8d "driven ... through
the actor shim"), over (a) the CPU oracle - the reference's own CPU-runnable case restated - and (b) libvsann.so on the GPU
with the vs_search coalescer on and off. Prints one JSON object.

  python scripts/c1_actor.py [--n 10000] [--nq 1000] [--clients 16] [--threads T]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import vector_store_text_b200 as vst  # noqa: E402


def run(lib, name, x, q, truth, k, threads, clients):
    n, d = x.shape
    nq = q.shape[0]
    a = vst.NativeActor("vector.ann", d, metric=vst.Metric.COS, capacity_hint=n, threads=threads, lib=lib)
    t0 = time.perf_counter()
    for i in range(n):
        a.add_or_replace(("pk", i), x[i])
    while a.count() != n:
        time.sleep(0.001)
    t_add = time.perf_counter() - t0
    a.ann(q[0], k)
    t0 = time.perf_counter()
    got = [a.ann(q[i], k)[0] for i in range(nq)]
    t_one = time.perf_counter() - t0
    out = [None] * nq

    def work(t):
        for i in range(t, nq, clients):
            out[i] = a.ann(q[i], k)[0]

    best = None
    for _ in range(3):
        th = [threading.Thread(target=work, args=(t,)) for t in range(clients)]
        t0 = time.perf_counter()
        [t.start() for t in th]
        [t.join() for t in th]
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    hits = sum(len({p[1] for p in got[i]} & set(truth[i].tolist())) for i in range(nq))
    st = a.stats()
    a.close()
    return {"backend": name, "actor_threads": st["threads"], "add_vectors_per_s": n / t_add, "add_seconds": t_add,
            "ann_qps_1_client": nq / t_one, "ann_latency_ms_1_client": 1e3 * t_one / nq,
            f"ann_qps_{clients}_clients": nq / best, "recall_at_10": hits / (nq * k), "max_in_flight": st["max_in_flight"], "max_queued": st["max_queued"]}


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--n", type=int, default=10000)
    p.add_argument("--nq", type=int, default=1000)
    p.add_argument("--clients", type=int, default=16)
    p.add_argument("--threads", type=int, default=0)
    p.add_argument("--only", default="", help="oracle | gpu | gpu-nocoalesce (internal)")
    a = p.parse_args()
    d, k = 128, 10
    x = np.random.default_rng(42).standard_normal((a.n, d), dtype=np.float32)
    q = np.random.default_rng(43).standard_normal((a.nq, d), dtype=np.float32)
    so = os.path.join(ROOT, "oracle", "liboracle_vsann.so")
    olib = vst.load_library(so, prefix="ovs_")
    o = vst.VsIndex(d, metric=vst.Metric.COS, capacity_hint=a.n, connectivity=2, expansion_add=8, lib=olib)
    o.set_threads(os.cpu_count() or 1)
    o.add_batch(np.arange(a.n, dtype=np.uint64), x)
    truth, _, _ = o.search(q, k, exact=True)
    o.close()
    if a.only:
        lib = olib if a.only == "oracle" else vst.load_library()
        print(json.dumps(run(lib, a.only, x, q, truth, k, a.threads, a.clients)))
        return
    res = []
    for name, env in (("oracle", {}), ("gpu", {}), ("gpu-nocoalesce", {"VSANN_COALESCE": "0"})):
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--n", str(a.n), "--nq", str(a.nq), "--clients", str(a.clients),
                            "--threads", str(a.threads), "--only", name], env={**os.environ, **env}, capture_output=True, text=True)
        if r.returncode != 0:
            res.append({"backend": name, "error": r.stderr[-400:]})
        else:
            res.append(json.loads(r.stdout.strip().splitlines()[-1]))
    print(json.dumps({"workload": f"configs[0]: {a.n} x {d} f32 cosine (i.i.d. normal, seeds 42/43), one add message per vector, {a.nq} ANN queries k={k}, "
                                  f"via the C++ index actor (channel 10, 2 x T in flight)", "host_cores": os.cpu_count(), "results": res}))


if __name__ == "__main__":
    main()
