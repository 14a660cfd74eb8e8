"""Build-rate variance probe: the bench's 1M x 128 build, R times in one process (device seconds and wall each)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vector_store_text_b200 as vst
n, d, R = int(os.environ.get("N", 1_000_000)), int(os.environ.get("D", 128)), int(os.environ.get("R", 3))
x = bench.gen(n, d, 42, "clustered"); xd = torch.from_numpy(x).cuda(); torch.cuda.synchronize()
keys = np.arange(n, dtype=np.uint64)
for r in range(R):
    g = vst.VsIndex(d, metric=vst.Metric.L2SQ, connectivity=16, expansion_add=128, expansion_search=64, capacity_hint=n)
    g.set_stats(os.environ.get("STATS", "0") == "1")
    torch.cuda.synchronize(); t0 = time.perf_counter()
    g.add_batch_device(keys, xd.data_ptr(), n); torch.cuda.synchronize()
    w = time.perf_counter() - t0
    st = g.stats()
    print(f"build {r}: device {st['build_seconds']:.3f} s  wall {w:.3f} s  {n / st['build_seconds'] / 1e6:.3f} M vectors/s", flush=True)
    g.close()
