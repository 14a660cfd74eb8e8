"""Exact search through the tcgen05 path vs the CPU oracle (ids and distance bits), with timing."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import vector_store_text_b200 as vst
n, d, nq, k = int(os.environ.get("N", 100000)), int(os.environ.get("D", 128)), int(os.environ.get("NQ", 1024)), int(os.environ.get("K", 10))
metric = {"l2": vst.Metric.L2SQ, "cos": vst.Metric.COS, "ip": vst.Metric.IP}[os.environ.get("METRIC", "l2")]
rng = np.random.default_rng(1)
x = rng.standard_normal((n, d), dtype=np.float32); q = rng.standard_normal((nq, d), dtype=np.float32)
dup = int(os.environ.get("DUP", 0))
if dup:      # adversarial: every vector appears `dup` times (exact ties), near-duplicates 1 ulp apart, queries on top of rows
    base = x[: n // dup]; x = np.tile(base, (dup, 1))[:n].copy(); x[::3] = np.nextafter(x[::3], np.float32(np.inf)); q[: nq // 2] = x[rng.integers(0, n, nq // 2)]
g = vst.VsIndex(d, metric=metric, capacity_hint=n, expansion_add=16, connectivity=4)
t0 = time.time(); g.add_batch(np.arange(n, dtype=np.uint64), x); g.flush(); print("build s", round(time.time() - t0, 2), flush=True)
for rep in range(3):
    t0 = time.time(); gk, gd, _ = g.search(q, k, exact=True); dt = time.time() - t0
    print(f"exact search {dt*1e3:.1f} ms  -> {nq*n*d*2/dt/1e12:.1f} TFLOP/s e2e", flush=True)
st = g.stats(); print("tc queries", st["exact_tc_queries"], "second scan", st["exact_tc_second_scan"], "fp32 fallback", st["exact_tc_fallback"], flush=True)
if os.environ.get("CHECK", "1") == "1":
    oracle = vst.load_library(os.path.join(ROOT, "oracle", "liboracle_vsann.so"), prefix="ovs_")
    o = vst.VsIndex(d, metric=metric, capacity_hint=n, lib=oracle, expansion_add=8, connectivity=2); o.set_threads(os.cpu_count() or 1)
    # the oracle's exact search does not need its graph; cap the build cost with tiny M / ef
    o.add_batch(np.arange(n, dtype=np.uint64), x)
    ok, od, _ = o.search(q, k, exact=True)
    print("ids equal", np.array_equal(gk, ok), "dist bits equal", np.array_equal(gd.view(np.uint32), od.view(np.uint32)), "mismatching queries", int((gk != ok).any(axis=1).sum()))
