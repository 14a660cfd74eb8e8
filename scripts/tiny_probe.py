"""Smallest end-to-end run of the hot path (used under compute-sanitizer)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vector_store_text_b200 as vst
n, d = int(os.environ.get("N", 3000)), int(os.environ.get("D", 128))
rng = np.random.default_rng(0)
x = rng.standard_normal((n, d), dtype=np.float32); q = rng.standard_normal((64, d), dtype=np.float32)
g = vst.VsIndex(d, metric=vst.Metric.L2SQ, capacity_hint=n)
g.add_batch(np.arange(n, dtype=np.uint64), x)
print("count", g.count())
a, _, _ = g.search(q, 10, ef=32)
e, _, _ = g.search(q, 10, exact=True)
print("recall", sum(len(set(r.tolist()) & set(t.tolist())) for r, t in zip(a, e)) / e.size)
