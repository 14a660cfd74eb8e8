"""Latency of vs_search_batch with host buffers for small batches on a configs[0]-sized index (10k x 128 cosine): what a round of the
index actor costs when it hands the library its whole in-flight window."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vector_store_text_b200 as vst
n, d, k = 10000, 128, 10
rng = np.random.default_rng(3)
x = rng.standard_normal((n, d), dtype=np.float32)
g = vst.VsIndex(d, metric=vst.Metric.COS, capacity_hint=n)
g.add_batch(np.arange(n, dtype=np.uint64), x); g.flush()
for nq in (1, 8, 16, 32, 64, 128):
    q = rng.standard_normal((nq, d), dtype=np.float32)
    for _ in range(20): g.search(q, k)
    t0 = time.perf_counter(); reps = 300
    for _ in range(reps): g.search(q, k)
    dt = (time.perf_counter() - t0) / reps
    print(f"nq={nq}: {dt * 1e3:.3f} ms per call -> {nq / dt / 1e3:.1f} k QPS if calls follow one another", flush=True)
