"""K3/K4 at other shapes (configs[2]: 768-d cosine, configs[4]: 1536-d dot): build rate, QPS, recall, HBM fraction by evaluation count."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vector_store_text_b200 as vst
n, d, nq = int(os.environ.get("N", 200000)), int(os.environ.get("D", 768)), int(os.environ.get("NQ", 4096))
metric = int(os.environ.get("METRIC", 0)); k = int(os.environ.get("K", 10)); efs = [int(e) for e in os.environ.get("EFS", "32,64,128").split(",")]
x = bench.gen(n, d, 42, "clustered"); q = bench.gen(nq, d, 43, "clustered")
xd = torch.from_numpy(x).cuda(); qd = torch.from_numpy(q).cuda()
g = vst.VsIndex(d, metric=vst.Metric(metric), connectivity=16, expansion_add=128, expansion_search=64, capacity_hint=n)
g.add_batch_device(np.arange(n, dtype=np.uint64), xd.data_ptr(), n)
st = g.stats(); print(f"n={n} d={d} metric={metric}: build {n / st['build_seconds'] / 1e3:.0f} k vectors/s ({st['build_seconds']:.2f} s)", flush=True)
ok = torch.empty((nq, k), dtype=torch.int64, device="cuda"); od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), exact=True, stream=s); torch.cuda.synchronize()
truth = ok.cpu().numpy().copy()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for ef in efs:
    g.set_stats(True); g.reset_stats()
    g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), ef=ef, stream=s); torch.cuda.synchronize()
    sv = g.stats(); g.set_stats(False)
    got = ok.cpu().numpy()
    rec = np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(got, truth)])
    tot = 0.0
    for _ in range(5):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), ef=ef, stream=s); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / 5
    ev = sv["search_dist_evals"] / max(1, sv["search_queries"])
    gbs = ev * d * 4 * nq / (ms * 1e-3) / 1e9
    print(f"  ef={ef}: recall {rec:.4f}  {nq / ms * 1e3 / 1e6:.3f} M QPS  {ms:.3f} ms  evals/query {ev:.0f}  gathered {gbs:.0f} GB/s", flush=True)
