"""BASELINE.json configs[2..4] at their per-GPU shapes on ONE B200 (the bench line stays configs[1]; these are
recorded under profiles/ as evidence for the other shapes):

  c3  10M x 768 f32 cosine over 8 shards  -> one shard = 1.25M x 768, queries in batches of 1024, k=10
  c4  100M x 96 f32 L2sq at 1/2/4/8 GPUs  -> one shard of the 8-GPU split = 12.5M x 96 (--n 100000000 = the 1-GPU case)
  c5  5M x 1536 f32 dot product, k=100, query batches 1..4096, brute-force exact vs HNSW ef sweep (fits one GPU whole)

Rows are generated on the device (torch Philox generator, seed 42 corpus / 43 queries; clustered = 4096 Gaussian
centres + 0.3 sigma noise, rows L2-normalised for cos / ip) so that no 30 GB host array is needed. Build goes through
vs_add_batch_device, ground truth through the exact path (K1/K2), ANN through K3; every time is CUDA events on the
launching stream with a 256 MB L2 flush between repetitions. One JSON object per config on stdout.

  python scripts/config_sweep.py --config c5 [--n N] [--nq Q] [--dist clustered|gaussian]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vector_store_text_b200 as vst  # noqa: E402

CONFIGS = {
    "c3": dict(n=1_250_000, d=768, metric="COS", k=10, nq=10_240, batches=[1024], efs=[16, 24, 32, 48, 64, 96, 128],
               name="configs[2] 10M x 768 f32 cosine / 8 shards: one shard"),
    "c4": dict(n=12_500_000, d=96, metric="L2SQ", k=10, nq=10_000, batches=[10_000], efs=[16, 24, 32, 48, 64, 96, 128],
               name="configs[3] 100M x 96 f32 L2sq: one shard of the 8-GPU split"),
    "c5": dict(n=5_000_000, d=1536, metric="IP", k=100, nq=4096, batches=[1, 8, 64, 512, 4096], efs=[100, 128, 256, 512, 1024],
               name="configs[4] 5M x 1536 f32 dot product, k=100"),
}


def gen_device(n, d, seed, dist, normalise, dev, chunk=1 << 20):
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    out = torch.empty((n, d), dtype=torch.float32, device=dev)
    centres = None
    if dist == "clustered":
        gc = torch.Generator(device=dev)
        gc.manual_seed(4242)
        centres = torch.randn((4096, d), generator=gc, device=dev, dtype=torch.float32)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        blk = torch.randn((b - a, d), generator=g, device=dev, dtype=torch.float32)
        if centres is not None:
            idx = torch.randint(0, 4096, (b - a,), generator=g, device=dev)
            blk.mul_(0.3).add_(centres[idx])
        if normalise:
            blk.div_(blk.norm(dim=1, keepdim=True).clamp_min_(1e-30))
        out[a:b] = blk
    return out


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--config", required=True, choices=sorted(CONFIGS))
    p.add_argument("--n", type=int, default=0)
    p.add_argument("--nq", type=int, default=0)
    p.add_argument("--dist", default="clustered", choices=["clustered", "gaussian"])
    p.add_argument("--reps", type=int, default=5)
    p.add_argument("--M", type=int, default=16)
    p.add_argument("--ef-add", type=int, default=128)
    p.add_argument("--only-exact", action="store_true", help="brute-force legs only (use --M 2 --ef-add 8 for a cheap build)")
    p.add_argument("--batches", default="", help="comma-separated query batch sizes overriding the config's")
    a = p.parse_args()
    c = dict(CONFIGS[a.config])
    n, nq, d, k = a.n or c["n"], a.nq or c["nq"], c["d"], c["k"]
    if a.batches:
        c["batches"] = [int(b) for b in a.batches.split(",")]
    metric = getattr(vst.Metric, c["metric"])
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6545.6))
    tf32 = float(peaks.get("bf16_tflops", 1672.9)) / 2

    t0 = time.perf_counter()
    norm = c["metric"] in ("COS", "IP")
    xd = gen_device(n, d, 42, a.dist, norm, dev)
    qd = gen_device(nq, d, 43, a.dist, norm, dev)
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t0

    g = vst.VsIndex(d, metric=metric, connectivity=a.M, expansion_add=a.ef_add, expansion_search=64, capacity_hint=n, device=0)
    keys = np.arange(n, dtype=np.uint64)
    t0 = time.perf_counter()
    g.add_batch_device(keys, xd.data_ptr(), n)
    torch.cuda.synchronize()
    build_wall = time.perf_counter() - t0
    st = g.stats()
    assert g.count() == n
    out = {"config": a.config, "workload": f"{c['name']}: {n} x {d} {c['metric']} ({a.dist}, device Philox seed 42/43), M={a.M} ef_add={a.ef_add}, "
                                           f"{nq} queries k={k}",
           "generate_s": t_gen,
           "build": {"vectors_per_s": n / st["build_seconds"], "seconds_device": st["build_seconds"], "seconds_wall": build_wall,
                     "batches": st["build_batches"], "device_bytes": st["device_bytes"]}}
    print(f"# built {n} x {d}: {n / st['build_seconds'] / 1e3:.0f} k vectors/s ({st['build_seconds']:.2f} s device, {build_wall:.2f} s wall)",
          file=sys.stderr, flush=True)
    del xd          # the index holds its own copy; free the source rows
    torch.cuda.empty_cache()

    ok = torch.empty((nq, k), dtype=torch.int64, device=dev)
    od = torch.empty((nq, k), dtype=torch.float32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def run(b, ef=0, exact=False, off=0):
        g.search_device(b, qd.data_ptr() + off * d * 4, k, ok.data_ptr() + off * k * 8, od.data_ptr() + off * k * 4, ef=ef, exact=exact, stream=s)

    def timed(b, ef=0, exact=False, reps=a.reps):
        """mean ms of one batch of b queries (CUDA events, L2 flushed before each repetition)."""
        run(b, ef, exact)
        torch.cuda.synchronize()
        tot = 0.0
        for r in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run(b, ef, exact, off=((r * b) % max(1, nq - b + 1)) if b < nq else 0)
            e1.record()
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / reps

    # exact ground truth (also the brute-force leg)
    run(nq, exact=True)
    torch.cuda.synchronize()
    truth = ok.cpu().numpy().copy()
    sx = g.stats()
    out["exact"] = {"tc_queries": sx["exact_tc_queries"], "second_scan_queries": sx["exact_tc_second_scan"],
                    "fp32_fallback_queries": sx["exact_tc_fallback"], "by_batch": {}}
    g.reset_stats()
    for b in sorted(set(c["batches"] + [nq])):
        if b > nq:
            continue
        s0 = g.stats()
        ms = timed(b, exact=True)
        s1 = g.stats()
        fl, by = 2.0 * b * n * d, n * d * 4.0
        path = {kk: s1[kk] - s0[kk] for kk in ("exact_tc_queries", "exact_tc_second_scan", "exact_tc_fallback", "exact_stream_queries")}
        out["exact"]["by_batch"][str(b)] = {"ms": ms, "qps": b / ms * 1e3, "tflops": fl / ms / 1e9, "frac_tf32_peak": fl / ms / 1e9 / tf32,
                                           "corpus_gbs": by / ms / 1e6, "frac_hbm_peak": by / ms / 1e6 / hbm, "queries_by_path": path}
        print(f"# exact B={b}: {ms:.3f} ms  {b / ms * 1e3:.0f} QPS  {fl / ms / 1e9:.0f} TFLOP/s  {by / ms / 1e6:.0f} GB/s   "
              f"[stream {path['exact_stream_queries']} tc {path['exact_tc_queries']} 2nd-scan {path['exact_tc_second_scan']} fp32 {path['exact_tc_fallback']}]",
              file=sys.stderr, flush=True)
    out["exact"]["stream_queries"] = g.stats().get("exact_stream_queries", 0)
    if a.only_exact:
        out["peaks"] = {"hbm_gbs": hbm, "tf32_tflops": tf32}
        g.close()
        print(json.dumps(out), flush=True)
        return

    # ANN: recall@k and QPS by ef on the whole query set, then by batch size at the smallest ef with recall >= 0.95
    out["ann"] = {"sweep": {}, "by_batch": {}}
    chosen = None
    for ef in c["efs"]:
        if ef < k:
            continue
        g.set_stats(True)
        g.reset_stats()
        run(nq, ef)
        torch.cuda.synchronize()
        sv = g.stats()
        g.set_stats(False)
        got = ok.cpu().numpy()
        rec = float(np.mean([len(np.intersect1d(x, y)) / k for x, y in zip(got, truth)]))
        ms = timed(nq, ef)
        ev = sv["search_dist_evals"] / max(1, sv["search_queries"])
        ex = sv["search_expansions"] / max(1, sv["search_queries"])
        gbs = (ev * d * 4 + ex * 2 * a.M * 4) * nq / ms / 1e6
        out["ann"]["sweep"][str(ef)] = {"recall": rec, "ms": ms, "qps": nq / ms * 1e3, "kernel_dist_evals_per_query": ev,
                                       "kernel_expansions_per_query": ex, "gathered_gbs": gbs, "frac_hbm_peak": gbs / hbm}
        print(f"# ann ef={ef}: recall@{k} {rec:.4f}  {nq / ms * 1e3 / 1e6:.3f} M QPS  {ms:.3f} ms  evals/query {ev:.0f}  {gbs:.0f} GB/s",
              file=sys.stderr, flush=True)
        if chosen is None and rec >= 0.95:
            chosen = ef
    out["ann"]["ef_at_recall95"] = chosen
    efb = chosen or c["efs"][-1]
    for b in c["batches"]:
        if b > nq:
            continue
        ms = timed(b, efb)
        out["ann"]["by_batch"][str(b)] = {"ef": efb, "ms": ms, "qps": b / ms * 1e3}
        print(f"# ann B={b} ef={efb}: {ms:.3f} ms  {b / ms * 1e3:.0f} QPS", file=sys.stderr, flush=True)
    out["peaks"] = {"hbm_gbs": hbm, "tf32_tflops": tf32, "source": "MEASURED_PEAKS.json (TF32 = bf16 burst / 2)" if peaks else "B200_PROFILING.md fallback"}
    g.close()
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
