#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json: "index build vectors/s; ANN QPS at
recall@10>=0.95, 1/2/4/8 B200 vs CPU").

Default workload = BASELINE configs[1] (`--config c2`): SIFT-1M-shaped synthetic, 1M x 128-d f32 L2, M=16,
ef_add=128, 10k queries, k=10 on one B200; N ranks => N shards of 1M vectors (weak scaling: the corpus grows
N x, every rank answers every query on its shard, the per-query top-k is merged over NVLink).
`--config c3|c4|c5` selects BASELINE configs[2..4] at FULL size (rows generated on the device, strong scaling:
the corpus is split over the ranks).

A step = one pass of the query batch through the ANN search (K3) and, at N > 1, the exchange + merge (K6), at
the smallest ef of the grid whose recall@k against exact ground truth is >= 0.95.
  value      queries answered per second against the WHOLE corpus, everything resident in HBM (CUDA events)
  e2e        the same batch through the C ABI with pinned HOST buffers (H2D + D2H inside the timed region)
  build      vectors/s on WALL time from the first vs_add_batch of host rows to vs_count() == N, its HBM roofline
  roofline   algorithmic bytes of the traversal (oracle beam on the GPU-built graph) / kernel time / measured HBM peak
  cpu_baseline / --impl reference   the CPU restatement (oracle/, SIMD distances) on the SAME corpus, queries, ef grid

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config c2|c3|c4|c5] [--dist clustered|gaussian]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

M, EF_ADD, EF_PARITY, RECALL_TARGET = 16, 128, 64, 0.95
EF_GRID = {10: [16, 18, 20, 22, 24, 28, 32, 40, 48, 56, 64, 80, 96, 128, 192, 256, 384, 512, 768, 1024],
           100: [100, 128, 192, 256, 384, 512, 768, 1024]}

# BASELINE.json configs[1..4]. scaling "weak": `n` vectors PER RANK; "strong": `n` vectors in total, split over the ranks.
CONFIGS = {
    "c2": dict(name="configs[1] SIFT-1M-shaped synthetic", n=1_000_000, d=128, metric="L2SQ", k=10, nq=10_000, scaling="weak", gen="host"),
    "c3": dict(name="configs[2] text-embedding shape", n=10_000_000, d=768, metric="COS", k=10, nq=10_240, scaling="strong", gen="device",
               batches=[1024]),
    "c4": dict(name="configs[3] Deep-100M shape", n=100_000_000, d=96, metric="L2SQ", k=10, nq=10_000, scaling="strong", gen="device"),
    "c5": dict(name="configs[4] high-QPS serving", n=5_000_000, d=1536, metric="IP", k=100, nq=4096, scaling="strong", gen="device",
               batches=[1, 8, 64, 512, 4096], recall_sweep=True),
}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    p.add_argument("--n", type=int, default=0, help="override the corpus size (per rank for c2, total otherwise)")
    p.add_argument("--nq", type=int, default=0)
    p.add_argument("--dist", default="clustered", choices=["clustered", "gaussian"])
    p.add_argument("--cpu-n", type=int, default=0, help="cap the CPU baseline's corpus (0 = the full corpus for c2, a stated subsample otherwise)")
    p.add_argument("--no-cpu", action="store_true", help="skip every CPU leg (profiling runs)")
    p.add_argument("--no-extra", action="store_true", help="skip the i.i.d. block, the exact-mode block and the via-actor leg")
    p.add_argument("--ef", type=int, default=0, help="skip the sweep and use this ef")
    p.add_argument("--capacity-hint", type=int, default=0, help="initial reserve / growth increment per shard (0 = the shard size)")
    p.add_argument("--one-build", action="store_true", help="profiling runs: only the host-row build (no warm-up build, no device-resident build)")
    p.add_argument("--single-process", action="store_true", help="N > 1 through ONE multi-device vs_index (vs_options.n_devices) instead of torchrun")
    return p.parse_args()


# ------------------------------------------------------------------------------------------ data
def gen(n, d, seed, dist, centres_seed=4242):
    """NumPy PCG64 standard_normal float32, corpus seed 42, query seed 43; the clustered
    variant draws 4096 Gaussian centres and adds 0.3-sigma noise (SIFT-like low intrinsic dimension)."""
    rng = np.random.default_rng(seed)
    if dist == "gaussian":
        return rng.standard_normal((n, d), dtype=np.float32)
    c = np.random.default_rng(centres_seed).standard_normal((4096, d), dtype=np.float32)
    a = rng.integers(0, 4096, n)
    x = rng.standard_normal((n, d), dtype=np.float32)
    x *= np.float32(0.3)
    x += c[a]
    return x


class DeviceRows:
    """Rows generated on the device, chunk by chunk (no 38 GB host array): torch's CUDA generator is
    Philox; seed = base + 1000 * shard, one generator per shard, chunks drawn in order so any chunking gives the same rows."""

    def __init__(self, torch, dev, d, seed, dist, normalise):
        self.torch, self.dev, self.d, self.normalise = torch, dev, d, normalise
        self.g = torch.Generator(device=dev)
        self.g.manual_seed(seed)
        self.centres = None
        if dist == "clustered":
            gc = torch.Generator(device=dev)
            gc.manual_seed(4242)
            self.centres = torch.randn((4096, d), generator=gc, device=dev, dtype=torch.float32)

    def next(self, rows):
        t = self.torch
        blk = t.randn((rows, self.d), generator=self.g, device=self.dev, dtype=t.float32)
        if self.centres is not None:
            idx = t.randint(0, 4096, (rows,), generator=self.g, device=self.dev)
            blk.mul_(0.3).add_(self.centres[idx])
        if self.normalise:
            blk.div_(blk.norm(dim=1, keepdim=True).clamp_min_(1e-30))
        return blk


def recall(found, truth):
    hit = 0
    for f, t in zip(found, truth):
        hit += len(np.intersect1d(f, t, assume_unique=False))
    return hit / truth.size


def shard_size(args, cfg, world):
    if cfg["scaling"] == "weak":
        return args.n or cfg["n"]
    return (args.n or cfg["n"]) // world


def workload_name(args, cfg, world):
    n = shard_size(args, cfg, world)
    lay = f"{n} x {cfg['d']} per shard x {world} shard(s)" if cfg["scaling"] == "weak" else f"{n * world} x {cfg['d']} over {world} shard(s)"
    return (f"{cfg['name']}: {lay}, f32 {cfg['metric']} ({args.dist}), M={M} ef_add={EF_ADD}, {args.nq or cfg['nq']} queries k={cfg['k']}")


def config_dict(args, cfg, world):
    """Identical in both arms (it is a function of the command line only); what a run measured goes to `operating_point`."""
    n = shard_size(args, cfg, world)
    return {"workload": workload_name(args, cfg, world), "config": args.config, "corpus_total": n * world, "dims": cfg["d"], "metric": cfg["metric"],
            "M": M, "ef_add": EF_ADD, "k": cfg["k"], "queries": args.nq or cfg["nq"], "recall_target": RECALL_TARGET,
            "ef_search": "smallest ef of ef_grid with recall@k >= recall_target on the arm's own index (see operating_point)",
            "ef_grid": EF_GRID[cfg["k"]], "ef_parity_point": EF_PARITY, "dist": args.dist,
            "generator": "numpy PCG64 seeds 42(+1000 r)/43" if cfg["gen"] == "host" else "torch CUDA Philox seeds 42(+1000 r)/43, rows generated on the device",
            "l2": "corpus + links per shard exceed the 126 MB L2; a 256 MB buffer is rewritten between timed steps",
            "parallelism": f"shard-by-key x{world}"}


# ------------------------------------------------------------------------------------------ clocks
_POLLER = r"""
import os, subprocess, sys, time
gpu, uuid, out = int(sys.argv[1]), sys.argv[2], open(sys.argv[3], "w", buffering=1)
try:   # stay out of the timed process's way: lowest priority, last core
    os.nice(19)
    cpus = sorted(os.sched_getaffinity(0))
    if len(cpus) > 2:
        os.sched_setaffinity(0, {cpus[-1]})
except Exception:
    pass
Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
try:
    import pynvml as nv
    nv.nvmlInit()
    h = None
    for u in ([uuid, "GPU-" + uuid] if uuid else []):
        try:
            h = nv.nvmlDeviceGetHandleByUUID(u); break
        except Exception:
            h = None
    if h is None:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        tok = vis.split(",")[gpu].strip() if vis else str(gpu)
        h = nv.nvmlDeviceGetHandleByIndex(int(tok)) if tok.isdigit() else nv.nvmlDeviceGetHandleByUUID(tok)
    mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
    reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
    masks = (0x8, 0x40, 0x20, 0x4)   # hw_slowdown, hw_thermal_slowdown, sw_thermal_slowdown, sw_power_cap (nvml.h)
    out.write("# nvml %d\n" % mx)
    while True:
        mhz = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        try:
            b = int(reasons(h))
        except Exception:
            b = 0
        out.write("%.6f,%d,%d,%d,%d,%d\n" % ((time.time(), mhz) + tuple(1 if b & m else 0 for m in masks)))
        time.sleep(0.002)
except Exception as e:
    out.write("# nvidia-smi (nvml failed: %s)\n" % (str(e)[:80],))
    p = subprocess.Popen(["nvidia-smi", "-i", str(gpu), "--query-gpu=" + Q, "--format=csv,noheader,nounits", "-lms", "20"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    for line in p.stdout:
        c = [v.strip() for v in line.split(",")]
        if len(c) >= 6 and c[0].replace(".", "").isdigit():
            out.write("%.6f,%s,%s,%s,%s,%s,%s\n" % ((time.time(), c[0]) + tuple("1" if v.lower().startswith("active") else "0" for v in c[2:6]) + (c[1],)))
"""


class ClockSampler:
    """SM clocks + throttle reasons under load (B200_PROFILING.md clocks line). The timed regions are a few
    milliseconds long, shorter than one `nvidia-smi -lms` period, so a separate process (no GIL shared with the
    timed loop) polls NVML every 2 ms (niced, pinned to the last core) and timestamps each sample; `mark(True/False)` records the wall-clock
    windows of the timed regions and the medians are taken over the samples that fell inside them. When NVML cannot
    be loaded the poller falls back to `nvidia-smi -lms 20`, and bench.py keeps the timed step itself running until
    a few samples have arrived (`need_tail`)."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu, uuid=None):
        self.gpu, self.uuid, self.proc, self.path = gpu, uuid or "", None, None
        self.windows, self.t_open = [], None

    def start(self):
        try:
            import tempfile
            fd, self.path = tempfile.mkstemp(prefix="vsann_clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen([sys.executable, "-c", _POLLER, str(self.gpu), self.uuid, self.path],
                                         stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def mark(self, inside):
        if inside:
            self.t_open = time.time()
        elif self.t_open is not None:
            self.windows.append((self.t_open, time.time()))
            self.t_open = None

    def _parse(self):
        rows, source, mx = [], None, None
        try:
            for line in open(self.path):
                if line.startswith("#"):
                    tok = line[1:].split()
                    source = tok[0]
                    if source == "nvml" and len(tok) > 1:
                        mx = float(tok[1])
                    continue
                c = line.strip().split(",")
                if len(c) < 6:
                    continue
                t, mhz = float(c[0]), float(c[1])
                if len(c) > 6 and c[6].replace(".", "").isdigit():
                    mx = max(mx or 0.0, float(c[6]))
                inside = any(a <= t <= b for a, b in self.windows) or (self.t_open is not None and t >= self.t_open)
                rows.append((mhz, [self.NAMES[i] for i in range(4) if c[2 + i] == "1"], inside))
        except Exception:
            pass
        return rows, source, mx

    def need_tail(self, want=3):
        """True while fewer than `want` samples fell inside a timed region (only the nvidia-smi fallback is that slow)."""
        if self.proc is None:
            return False
        return sum(1 for r in self._parse()[0] if r[2]) < want

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock poller could not be started"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows, source, mx = self._parse()
        try:
            os.unlink(self.path)
        except Exception:
            pass
        ins = [r for r in rows if r[2]]
        use = ins if ins else rows
        sm = [r[0] for r in use]
        reasons = sorted({n for r in use for n in r[1]})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None, "sm_max_mhz": mx, "reasons": reasons,
                "samples": len(sm), "samples_inside_timed_regions": len(ins), "samples_total": len(rows), "source": source}


# ------------------------------------------------------------------------------------------ CPU legs (oracle/: the checker, timed as the baseline)
def load_oracle(vst):
    so = os.path.join(ROOT, "oracle", "liboracle_vsann.so")
    if not os.path.exists(so):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    return vst.load_library(so, prefix="ovs_")


def host_cores():
    return int(os.environ.get("SCYLLA_USEARCH_BACKGROUND_THREADS", "0")) or (os.cpu_count() or 1)   # README.md:14-15


def numpy_truth(x, keys, q, k, metric):
    """Exact top-k ids on the host (BLAS), used where no GPU may be touched (--impl reference). f32 matmul is precise
    enough to rank neighbours for a recall figure (it is NOT the bit-exact oracle definition)."""
    out = np.empty((q.shape[0], k), dtype=np.uint64)
    xn = (x * x).sum(1) if metric == "L2SQ" else None
    step = int(max(8, min(256, (1 << 28) // max(1, x.shape[0]))))     # <= 1 GiB of scores per block of queries
    for a in range(0, q.shape[0], step):
        s = q[a:a + step] @ x.T
        if metric == "L2SQ":
            s = xn[None, :] - 2.0 * s
        elif metric == "COS":
            s = -s / np.maximum(np.linalg.norm(x, axis=1)[None, :], 1e-30)
        else:
            s = -s
        idx = np.argpartition(s, k, axis=1)[:, :k]
        part = np.take_along_axis(s, idx, 1)
        order = np.argsort(part, axis=1)
        out[a:a + step] = keys[np.take_along_axis(idx, order, 1)]
    return out


class CpuIndex:
    """The CPU restatement at the workload's parameters, built with all host threads (direct library calls = drive mode (a)
    of BASELINE.md §3), SIMD distances on (what usearch's SimSIMD kernels do per distance)."""

    def __init__(self, vst, d, metric, k, n_hint):
        self.vst, self.k, self.cores = vst, k, host_cores()
        self.o = vst.VsIndex(d, metric=getattr(vst.Metric, metric), connectivity=M, expansion_add=EF_ADD, expansion_search=EF_PARITY,
                             capacity_hint=n_hint, lib=load_oracle(vst))
        self.o.set_threads(self.cores)
        self.simd = True
        try:
            self.o.set_simd(True)
        except Exception:
            self.simd = False
        self.n, self.build_s = 0, 0.0

    def add(self, keys, x):
        t0 = time.perf_counter()
        self.o.add_batch(keys, x)
        self.build_s += time.perf_counter() - t0
        self.n += len(keys)

    def sweep(self, q, truth, grid, stop_at_parity=True):
        """recall@k by ef on (a sample of) the queries; returns (table, smallest ef reaching the target or None)."""
        tab, pick = {}, None
        for ef in grid:
            ak, _, _ = self.o.search(q, self.k, ef=ef)
            r = recall(ak, truth)
            tab[ef] = round(r, 4)
            if pick is None and r >= RECALL_TARGET:
                pick = ef
            if pick is not None and (not stop_at_parity or ef >= EF_PARITY):
                break
        return tab, pick

    def time_steps(self, q, ef, steps, warmup, budget_s=60.0):
        for _ in range(warmup):
            self.o.search(q, self.k, ef=ef)
        ts = []
        for _ in range(max(1, steps)):
            t0 = time.perf_counter()
            self.o.search(q, self.k, ef=ef)
            ts.append(time.perf_counter() - t0)
            if sum(ts) > budget_s:
                break
        return ts

    def work(self, q, ef):
        s0 = self.o.stats()
        self.o.search(q, self.k, ef=ef)
        s1 = self.o.stats()
        qn = max(1, s1["search_queries"] - s0["search_queries"])
        return (s1["search_dist_evals"] - s0["search_dist_evals"]) / qn, (s1["search_expansions"] - s0["search_expansions"]) / qn

    def close(self):
        self.o.close()


def via_actor_leg(vst, with_gpu):
    """Drive mode (b) of BASELINE.md §3 on configs[0] (the reference's own CPU-runnable case): 10k x 128 cosine, one AddOrReplace
    message per vector and one Ann message per query through libvsactor.so (channel 10, 2 x T in flight, usearch.rs:102,118-131)
    with native client threads (lib/actor_load), over the CPU restatement and, beside it, over libvsann.so."""
    exe = os.path.join(ROOT, "vector-store-text_b200", "lib", "actor_load")
    if not os.path.exists(exe):
        return {"unavailable": "lib/actor_load is not built"}
    out = {"config": "configs[0]: 10k x 128 f32 cosine, add + 1k ANN queries k=10 via the index actor, T = host cores, 2T clients"}
    T = host_cores()
    legs = [("cpu", os.path.join(ROOT, "oracle", "liboracle_vsann.so"), "ovs_")]
    if with_gpu:
        legs.append(("gpu", vst.lib_path(), "vs_"))
    for name, so, pre in legs:
        try:
            env = dict(os.environ, OVS_SIMD="1")
            r = subprocess.run([exe, so, pre, "10000", "128", "1000", "10", str(2 * T), str(T), "3"], capture_output=True, text=True, timeout=300, env=env)
            out[name] = json.loads(r.stdout.strip().splitlines()[-1])
        except Exception as e:   # noqa: BLE001
            out[name] = {"failed": f"{type(e).__name__}: {str(e)[:160]}"}
    return out


# ------------------------------------------------------------------------------------------ --impl reference
def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores. The reference's own code cannot be
    built here (usearch crate absent and unpinned, no Rust toolchain: DESIGN.md §2), so this times the oracle port - on the SAME
    corpus (all `corpus_total` vectors for c2), the same query batch, the same ef grid, K timed steps after W warm-ups."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import vector_store_text_b200 as vst
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    cfg = CONFIGS[args.config]
    d, k, metric = cfg["d"], cfg["k"], cfg["metric"]
    nq = args.nq or cfg["nq"]
    n_shard = shard_size(args, cfg, world)
    total = n_shard * world
    # configs[1]: the whole corpus up to 4M vectors (N <= 4 shards); beyond that a stated 4M-vector part, so that the arm still ends within minutes
    cpu_n = args.cpu_n or (min(total, 4_000_000) if cfg["gen"] == "host" else min(total, 200_000))
    cpu = CpuIndex(vst, d, metric, k, cpu_n)
    xs, ks = [], []
    if cfg["gen"] == "host":
        per = -(-cpu_n // world)
        for r in range(world):   # the GPU arm's shards: rows of shard r = gen(seed 42 + 1000 r), keys i * world + r
            m = min(per, n_shard, cpu_n - sum(len(a) for a in ks))
            if m <= 0:
                break
            x = gen(n_shard, d, 42 + 1000 * r, args.dist)[:m]
            keys = np.arange(m, dtype=np.uint64) * world + r
            cpu.add(keys, x)
            xs.append(x); ks.append(keys)
        q = gen(nq, d, 43, args.dist)
        sample = (f"oracle HNSW (M={M}, ef_add={EF_ADD}, AVX2 distances={cpu.simd}) built on {cpu.n} of the {total} vectors of the workload "
                  f"with {cpu.cores} threads; every step = the workload's {nq} queries, k={k}")
    else:   # device-generated shapes: the same distribution from the host generator, a stated subsample
        x = gen(cpu_n, d, 42, args.dist)
        q = gen(nq, d, 43, args.dist)
        if metric in ("COS", "IP"):
            x /= np.linalg.norm(x, axis=1, keepdims=True); q /= np.linalg.norm(q, axis=1, keepdims=True)
        keys = np.arange(cpu_n, dtype=np.uint64)
        cpu.add(keys, x)
        xs.append(x); ks.append(keys)
        sample = (f"oracle HNSW (M={M}, ef_add={EF_ADD}, AVX2 distances={cpu.simd}) on a {cpu_n}-vector SUBSAMPLE of the workload's distribution "
                  f"({total} vectors would take hours on the host), {cpu.cores} threads; every step = {nq} queries, k={k}")
    xa, ka = np.concatenate(xs), np.concatenate(ks)
    ts_n = min(1000 if len(xa) <= 2_000_000 else 200, nq)
    truth = numpy_truth(xa, ka, q[:ts_n], k, metric)
    tab, pick = cpu.sweep(q[:ts_n], truth, [args.ef] if args.ef else EF_GRID[k])
    reached = pick is not None
    ef = pick or max(tab)
    ts = cpu.time_steps(q, ef, args.steps, args.warmup, budget_s=240.0)
    qps_by_ef = {}
    for e in [e for e in EF_GRID[k] if e <= EF_PARITY and e != ef][:12]:   # one step each: whatever ef the GPU arm settles on has its CPU number here
        t = cpu.time_steps(q, e, 1, 0)
        qps_by_ef[e] = nq / t[0]
    qps_by_ef[ef] = nq / float(np.mean(ts))
    ev, ex = cpu.work(q[:ts_n], ef)
    st = cpu.o.stats()
    val = nq / float(np.mean(ts))
    cb = {"value": val, "unit": "queries/s", "cores": cpu.cores, "kind": "port", "sample": sample,
          "build_vectors_per_s": cpu.n / cpu.build_s, "build_seconds": cpu.build_s, "build_dist_evals_per_vector": st["build_dist_evals"] / max(1, cpu.n),
          "recall_at_k": tab.get(ef), "ef_search": ef, "recall_by_ef": tab, "qps_by_ef": {str(e): v for e, v in sorted(qps_by_ef.items())},
          "dist_evals_per_query": ev, "expansions_per_query": ex,
          "recall_ground_truth": f"host BLAS exact top-{k} on the first {ts_n} queries (no GPU on this arm)"}
    line = {"impl": "reference", "metric": "ann_qps_at_recall95", "value": val, "unit": "queries/s", "n_gpus": args.gpus,
            "steps": len(ts), "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(ts)), "higher_is_better": True, "scaling": cfg["scaling"],
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args, cfg, world),
            "operating_point": {"ef_search": ef, "recall_at_k": tab.get(ef), "recall_target_reached": reached, "ef_sweep_recall": tab},
            "build": {"vectors_per_s": cpu.n / cpu.build_s, "seconds_wall": cpu.build_s, "vectors": cpu.n, "unit": "vectors/s"},
            "cpu_baseline": cb, "e2e": {"value": val, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    cpu.close()
    emit(line)


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout. Libraries print there too (NCCL's "NCCL version ..." banner goes to fd 1 at
    init), so fd 1 is pointed at stderr for the duration of the run and the line is written to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def log(msg):
    print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------ the GPU arm
def algorithmic_counts(vst, g, x, q, ef, k, sample=200):
    """Work of the sequential algorithm (oracle beam, oracle/oracle_vsann.cpp ovs_graph_search) on the
    GPU-built graph for a query sample: distance evaluations and expansions per query."""
    olib = load_oracle(vst)
    d = x.shape[1]
    links0, _ = g.export_level(0)
    n = x.shape[0]
    uppers = []
    lvl = 1
    while M ** lvl < n:
        uppers.append(g.export_level(lvl)[0])
        lvl += 1
    ev_tot, ex_tot = 0, 0
    slots = np.zeros(k, np.uint32); dist = np.zeros(k, np.float32)
    for i in range(min(sample, q.shape[0])):
        qi = q[i]
        cur, dcur = 0, float(((x[0] - qi) ** 2).sum())
        ev, ex = 1, 0
        for l in range(len(uppers), 0, -1):          # greedy descent on the upper levels
            tab, stride = uppers[l - 1], M ** l
            while True:
                row = tab[cur // stride]; row = row[row != 0xFFFFFFFF]
                ex += 1
                if row.size == 0:
                    break
                dd = ((x[row] - qi) ** 2).sum(1); ev += row.size
                j = int(dd.argmin())
                if dd[j] < dcur:
                    cur, dcur = int(row[j]), float(dd[j])
                else:
                    break
        found, e1, e2 = C.c_size_t(), C.c_uint64(), C.c_uint64()
        olib.ovs_graph_search(1, d, n, x.ctypes.data_as(C.c_void_p), links0.ctypes.data_as(C.c_void_p), links0.shape[1], None, cur,
                              qi.ctypes.data_as(C.c_void_p), k, ef, slots.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p),
                              C.byref(found), C.byref(e1), C.byref(e2))
        ev_tot += ev + e1.value; ex_tot += ex + e2.value
    s = min(sample, q.shape[0])
    return ev_tot / s, ex_tot / s


def main():
    args = parse()
    quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import vector_store_text_b200 as vst

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = vst.load_library()          # the CUDA library is mandatory; raises if missing
    try:
        uuid = str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        uuid = None
    sampler = ClockSampler(local, uuid)   # started now (the poller needs ~1 s to come up), read after the e2e region
    if rank == 0:
        sampler.start()

    cfg = CONFIGS[args.config]
    d, k, metric_name = cfg["d"], cfg["k"], cfg["metric"]
    metric = getattr(vst.Metric, metric_name)
    nq = args.nq or cfg["nq"]
    # --single-process: ONE vs_index over args.gpus devices inside this process (vs_options.n_devices); same shards, same keys as torchrun
    G = args.gpus if (args.single_process and world == 1 and args.gpus > 1) else 1
    S = world * G                                   # shards behind the workload
    n = shard_size(args, cfg, S)
    n_local = n * G                                 # rows this process adds
    host_gen = cfg["gen"] == "host"
    grid = [args.ef] if args.ef else EF_GRID[k]
    # Every launch of the timed regions (flush write, events, search, merge, copies) goes to ONE explicit stream: the library
    # enqueues on the handle it is given (a NULL handle would select its own stream, unordered with torch's).
    ts = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(ts)
    stream = ts.cuda_stream
    assert stream != 0

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce(v, op):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=op)
        torch.cuda.synchronize()
        return float(t.item())

    def max_over_ranks(v):
        return reduce(v, dist.ReduceOp.MAX) if world > 1 else v

    def new_index(hint):
        return vst.VsIndex(d, metric=metric, connectivity=M, expansion_add=EF_ADD, expansion_search=EF_PARITY,
                           capacity_hint=hint * G, device=local, lib=lib, devices=list(range(G)) if G > 1 else None)

    # ---- data: shard r holds keys i * world + r (shard = key % world) -------------
    keys = np.arange(n, dtype=np.uint64) * world + rank if G == 1 else np.arange(n_local, dtype=np.uint64)
    hint = args.capacity_hint or n
    build = {"unit": "vectors/s"}
    x = None
    if host_gen:
        if G == 1:
            x = gen(n, d, 42 + 1000 * rank, args.dist)
        else:                                       # key i * G + r = row i of shard r, as under torchrun
            x = np.empty((n_local, d), dtype=np.float32)
            for r in range(G):
                x[r::G] = gen(n, d, 42 + 1000 * r, args.dist)
        q = gen(nq, d, 43, args.dist)
        xd = torch.from_numpy(x).to(dev)
        torch.cuda.synchronize()
        # (1) untimed warm-up build (module load, allocator), (2) device-resident rows timed, (3) HOST rows through vs_add_batch on
        # the wall clock from the first add to count() == N - the index the searches below run on.
        if args.warmup > 0 and not args.one_build:
            w = new_index(hint); w.add_batch_device(keys, xd.data_ptr(), n_local); w.close()
        w = new_index(hint)
        barrier()
        t0 = time.perf_counter()
        if not args.one_build:
            w.add_batch_device(keys, xd.data_ptr(), n_local)
            assert w.count() == n_local
        dev_wall = max_over_ranks(time.perf_counter() - t0)
        st = w.stats()
        if not args.one_build:
            build["device_resident"] = {"vectors_per_s": n * S / max_over_ranks(dev_wall if G > 1 else st["build_seconds"]), "seconds_device": max_over_ranks(st["build_seconds"]),
                                    "seconds_wall": dev_wall, "rows": "already in HBM (vs_add_batch_device)"}
        w.close()
        del xd
        xh = torch.from_numpy(x).pin_memory()
        g = new_index(hint)
        barrier()
        t0 = time.perf_counter()
        step_rows = 1 << 18                     # the caller hands over 262 144 rows per call, as an ingest loop would
        step_rows *= G
        for a in range(0, n_local, step_rows):
            b = min(n_local, a + step_rows)
            g.add_batch(keys[a:b], xh[a:b].numpy())
        assert g.count() == n_local
        build_wall = max_over_ranks(time.perf_counter() - t0)
        st = g.stats()
        build.update({"vectors_per_s": n * S / build_wall, "seconds_wall": build_wall, "seconds_device": max_over_ranks(st["build_seconds"]),
                      "batches": st["build_batches"], "h2d_bytes": n_local * d * 4 + n_local * 8,
                      "definition": "wall clock from the first vs_add_batch (pinned HOST rows, H2D inside) to vs_count() == N"})
        del xh
    else:
        q = None
        normalise = metric_name in ("COS", "IP")
        rows = DeviceRows(torch, dev, d, 42 + 1000 * rank, args.dist, normalise)
        qd_full = DeviceRows(torch, dev, d, 43, args.dist, normalise).next(nq)
        g = new_index(hint)
        chunk = min(n_local, 4_000_000)
        barrier()
        wall = 0.0
        for a in range(0, n_local, chunk):
            b = min(n_local, a + chunk)
            blk = rows.next(b - a)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            g.add_batch_device(keys[a:b], blk.data_ptr(), b - a)
            wall += time.perf_counter() - t0
            del blk
            if rank == 0 and (a // chunk) % 5 == 0:
                log(f"built {b} / {n_local} rows in this process, {b / wall / 1e6:.2f} M vectors/s so far")
        assert g.count() == n_local
        build_wall = max_over_ranks(wall)
        st = g.stats()
        build.update({"vectors_per_s": n * S / build_wall, "seconds_wall": build_wall, "seconds_device": max_over_ranks(st["build_seconds"]),
                      "batches": st["build_batches"],
                      "definition": "wall clock of the vs_add_batch_device calls (rows generated in HBM chunk by chunk; generation not timed) up to vs_count() == N"})
        q = qd_full.cpu().numpy()
    build["capacity_growths"] = st.get("reserve_calls")
    build["reserve_seconds"] = st.get("reserve_seconds")
    build["device_bytes"] = st.get("device_bytes")
    build["tables_grow_in_place"] = st.get("tables_grow_in_place")
    qd = torch.from_numpy(q).to(dev)
    torch.cuda.synchronize()
    if rank == 0:
        log(f"build done: {build['vectors_per_s'] / 1e6:.3f} M vectors/s wall")

    ok = torch.empty((nq, k), dtype=torch.int64, device=dev)
    od = torch.empty((nq, k), dtype=torch.float32, device=dev)
    gk = torch.empty((world, nq, k), dtype=torch.int64, device=dev) if world > 1 else None
    gd = torch.empty((world, nq, k), dtype=torch.float32, device=dev) if world > 1 else None
    mk = torch.empty((nq, k), dtype=torch.int64, device=dev)
    md = torch.empty((nq, k), dtype=torch.float32, device=dev)

    # ---- exchange step at N > 1. Default: fused gather + merge over peer memory - every rank's result lists
    # live in symmetric memory (mapped into every peer over NVLink / NVSwitch); K6 waits for the peers' per-step flags and
    # loads their lists directly: no all-gather, no gathered copy. VSANN_P2P_MERGE=0, or symmetric memory failing to come up,
    # selects NCCL all_gather + the dense merge kernel.
    p2p = None
    if world > 1 and os.environ.get("VSANN_P2P_MERGE", "1") != "0":
        try:
            from vector_store_text_b200.sharded import PeerExchange   # the package's exchange: symmetric memory + fused merge
            p2p = PeerExchange(nq, k, dev, lib=lib)
        except Exception as e:   # noqa: BLE001
            print(f"[bench] symmetric memory unavailable ({type(e).__name__}: {str(e)[:120]}); using NCCL all_gather", file=sys.stderr)
            p2p = None
    align_starts = os.environ.get("VSANN_BENCH_ALIGN", "1") != "0"
    exchange = (p2p.describe() if p2p else "nccl all_gather + merge kernel") if world > 1 else (
        f"one vs_index over {G} devices in one process: fan-out on {G} streams, merge kernel on device 0 over peer mappings" if G > 1 else "none")

    def local_search(ef, exact=False, nb=None, off=0):
        b = nb or nq
        g.search_device(b, qd.data_ptr() + off * d * 4, k, ok.data_ptr() + off * k * 8, od.data_ptr() + off * k * 4, ef=ef, exact=exact, stream=stream)
        return ok

    def search_step(ef, exact=False):
        """One pass of the query batch: shard search, then (N > 1) the exchange + merge kernel (K6)."""
        if world == 1:
            return local_search(ef, exact)
        if p2p:
            i = p2p.search(g, nq, qd.data_ptr(), ef=ef, exact=exact, stream=stream)
            p2p.merge(i, mk, md)
            return mk
        local_search(ef, exact)
        dist.all_gather_into_tensor(gk.view(-1), ok.view(-1))
        dist.all_gather_into_tensor(gd.view(-1), od.view(-1))
        vst.merge_topk_device(local, world, nq, k, gk.data_ptr(), gd.data_ptr(), mk.data_ptr(), md.data_ptr(), stream=stream, lib=lib)
        return mk

    def keys_of(t):
        torch.cuda.synchronize()
        return t.cpu().numpy().view(np.uint64).copy()

    truth = keys_of(search_step(0, exact=True))
    truth_local = keys_of(local_search(0, exact=True)) if world > 1 else truth

    # ---- operating point. The efficiency curve needs the SAME criterion at every N: the smallest ef that reaches the target on a
    # shard searched ALONE (max over ranks) - the sharded search at that ef can only be better. The smallest ef that reaches the
    # target on the merged result is reported beside it (`smallest_ef_global`).
    sweep_local, sweep = {}, {}
    chosen = None
    for ef in grid:
        try:
            r = recall(keys_of(local_search(ef)), truth_local)
        except vst.VsError as e:
            sweep_local[ef] = f"unsupported: {e.message}"
            break
        sweep_local[ef] = round(r, 4)
        ok_everywhere = reduce(1.0 if r >= RECALL_TARGET else 0.0, dist.ReduceOp.MIN) if world > 1 else (1.0 if r >= RECALL_TARGET else 0.0)
        if ok_everywhere >= 1.0 and chosen is None:
            chosen = ef
        if chosen is not None and ef >= EF_PARITY:
            break
    reached = chosen is not None
    if chosen is None:
        chosen = max(e for e, r in sweep_local.items() if not isinstance(r, str))
    ef_global = None
    if world > 1:
        for ef in [e for e in grid if e <= chosen]:
            r = recall(keys_of(search_step(ef)), truth)
            sweep[ef] = round(r, 4)
            if r >= RECALL_TARGET and ef_global is None:
                ef_global = ef
    else:
        sweep, ef_global = dict(sweep_local), chosen
    parity_ef = max(EF_PARITY, k)
    if parity_ef not in sweep:
        sweep[parity_ef] = round(recall(keys_of(search_step(parity_ef)), truth), 4)

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def timed(fn, steps, warmup, mark=True):
        for _ in range(warmup):
            fn()
        barrier()
        tot = 0.0
        if mark:
            sampler.mark(True)
        for _ in range(steps):
            flush_buf.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if world > 1:
                dist.barrier()
                if p2p is not None and align_starts:
                    # the ranks leave the NCCL barrier (a host wait) tens of microseconds apart, and a rank that starts early would count
                    # the others' lateness as exchange time; in service the ranks ARE aligned (every step ends in the merge's flag
                    # barrier). One device-side barrier kernel on the timed stream, BEFORE the first event, puts them back in step.
                    p2p.fhdl.barrier(channel=1)
            e0.record(ts)
            fn()
            e1.record(ts)
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        if mark:
            sampler.mark(False)
        barrier()
        return max_over_ranks(tot / steps)

    ms = timed(lambda: search_step(chosen), args.steps, args.warmup)
    ms_parity = timed(lambda: search_step(parity_ef), max(2, args.steps // 2), 1)
    ms_global_ef = timed(lambda: search_step(ef_global), max(2, args.steps // 2), 1) if (ef_global and ef_global != chosen) else ms
    # the search kernel alone (no exchange) for the roofline
    ms_kernel = timed(lambda: local_search(chosen), args.steps, 1)
    if rank == 0:
        log(f"search: ef={chosen} recall={sweep.get(chosen)} {ms:.4f} ms per step ({nq / ms / 1e3:.2f} M queries/s), kernel {ms_kernel:.4f} ms")

    # ---- recall sweep (configs[4]: "brute-force exact vs HNSW recall sweep"): every ef of the grid, recall + QPS + gathered bytes/s ----
    recall_sweep = None
    if cfg.get("recall_sweep") or os.environ.get("VSANN_BENCH_SWEEP"):
        recall_sweep = {}
        for ef in grid:
            try:
                g.set_stats(True); g.reset_stats()
                r = recall(keys_of(search_step(ef)), truth)
                sv = g.stats(); g.set_stats(False)
            except vst.VsError as e:
                recall_sweep[str(ef)] = f"unsupported: {e.message}"
                break
            t = timed(lambda: search_step(ef), max(3, args.steps // 4), 1, mark=False)
            evq = sv["search_dist_evals"] / max(1, sv["search_queries"]); exq = sv["search_expansions"] / max(1, sv["search_queries"])
            recall_sweep[str(ef)] = {"recall_at_k": round(r, 4), "ms_per_step": t, "qps": nq / t * 1e3, "kernel_dist_evals_per_query": evq,
                                     "gathered_gbs_per_gpu": (evq * d * 4 + exq * 2 * M * 4) * nq / t / 1e6}
            if rank == 0:
                log(f"sweep ef={ef}: recall {r:.4f} {nq / t / 1e3:.3f} M queries/s, gathered {recall_sweep[str(ef)]['gathered_gbs_per_gpu']:.0f} GB/s per GPU")

    by_batch = {}
    for b in cfg.get("batches", []):
        if b <= nq and world == 1:
            reps, tot = max(3, args.steps // 2), 0.0
            for r in range(reps):
                off = (r * b) % max(1, nq - b + 1) if b < nq else 0
                tot += timed(lambda: local_search(chosen, nb=b, off=off), 1, 1 if r == 0 else 0, mark=False)
            by_batch[str(b)] = {"ef": chosen, "ms": tot / reps, "qps": b / (tot / reps) * 1e3}

    # ---- brute-force mode (K1 on tcgen05 + exact re-rank): the same batch, timed like a step -----
    exact = None
    if not args.no_extra:
        g.set_stats(True); g.reset_stats()
        ms_exact = timed(lambda: search_step(0, exact=True), max(2, args.steps // 2), 1, mark=False)
        st_exact = g.stats()
        g.set_stats(False)
        exact_by_batch = {}
        for b in cfg.get("batches", []):
            if b <= nq and world == 1:
                t = timed(lambda: local_search(0, exact=True, nb=b), 3, 1, mark=False)
                exact_by_batch[str(b)] = {"ms": t, "qps": b / t * 1e3, "tflops": 2.0 * b * n * d / t / 1e9, "corpus_gbs": n * d * 4.0 / t / 1e6}

        # the tcgen05 TF32 rate of the same kernel on this GPU, measured instead of assumed: the scan without its epilogue arithmetic
        # (TMA + MMA: what a plain GEMM loop of this tile shape reaches) and without its loads as well (the MMA issue rate alone);
        # VSANN_TC_FLAGS is a measurement switch of exact_tc.cu - those two launches return right after the scan, their results are not used
        tf32_measured = {}
        if world == 1 and st_exact["exact_tc_scans"]:
            for name, fl in (("tma_mma_no_epilogue", 8), ("mma_only", 24)):
                os.environ["VSANN_TC_FLAGS"] = str(fl)
                try:
                    g.set_stats(True); g.reset_stats()
                    for _ in range(3):
                        local_search(0, exact=True)
                    torch.cuda.synchronize()
                    sv = g.stats()
                    if sv["exact_tc_scans"] and sv["exact_tc_scan_seconds"] > 0:
                        tf32_measured[name] = 2.0 * nq * n * d / (sv["exact_tc_scan_seconds"] / sv["exact_tc_scans"]) / 1e12
                finally:
                    os.environ.pop("VSANN_TC_FLAGS", None)
                    g.set_stats(False)

    # ---- e2e: same batch through the C ABI with pinned HOST buffers (H2D + D2H inside) --------------
    qh = torch.from_numpy(q).pin_memory()
    kh = torch.empty((nq, k), dtype=torch.int64).pin_memory()
    dh = torch.empty((nq, k), dtype=torch.float32).pin_memory()
    fh = np.zeros(nq, dtype=np.uintp)
    fn_host = getattr(lib, "vs_search_batch")

    def e2e_step(mode):
        if world > 1 and p2p:
            # sharded serving path: the traversal kernel reads the pinned HOST queries where they lie (vs_search_batch*_device accept any
            # device-accessible pointer; page-locked memory is one under UVA), the shard's lists go to symmetric memory (pushed to every
            # peer while the traversal runs), the fused barrier + merge follows on the same stream, and only the merged top-k goes back
            # to the host. (Measured and dropped: every rank uploading 1 / N of the block and gathering the rest over NVLink - three more
            # launches per step cost more than the host traffic they save: 0.70 against 0.62 ms at N = 2, equal at N = 8.)
            i = p2p.search(g, nq, qh.data_ptr(), ef=chosen, stream=stream)
            p2p.merge(i, mk, md)
            kh.copy_(mk, non_blocking=True); dh.copy_(md, non_blocking=True)
            ts.synchronize()
            return
        stt = fn_host(g._h, nq, C.c_void_p(qh.data_ptr()), k, chosen, 0, C.c_void_p(kh.data_ptr()), C.c_void_p(dh.data_ptr()), fh.ctypes.data_as(C.c_void_p))
        assert stt == 0
        if world > 1:   # NCCL fallback: results come back to the host per shard; merge on device as the serving rank would
            ok.copy_(kh, non_blocking=True); od.copy_(dh, non_blocking=True)
            dist.all_gather_into_tensor(gk.view(-1), ok.view(-1)); dist.all_gather_into_tensor(gd.view(-1), od.view(-1))
            vst.merge_topk_device(local, world, nq, k, gk.data_ptr(), gd.data_ptr(), mk.data_ptr(), md.data_ptr(), stream=stream, lib=lib)
            kh.copy_(mk, non_blocking=True); dh.copy_(md, non_blocking=True)
            ts.synchronize()

    import gc
    e2e_modes = ["in_place"] if (world > 1 and p2p) else ["c_abi_host_buffers"]
    e2e_by_mode = {}
    for mode in e2e_modes:
        for _ in range(max(1, args.warmup) + (10 if world > 1 else 0)):   # (N > 1: the first steps after the device-timed regions still see one-off costs)
            e2e_step(mode)
        barrier()
        gc.collect(); gc.disable()          # a collection inside a 6 ms region would be a tenth of it
        if mode == e2e_modes[0]:
            sampler.mark(True)
        per_step = []
        # at least 100 steps when a step is short: single steps of 1-8 ms (host scheduling) turn up every few dozen steps at N > 1
        # and would be a tenth of a 20-step mean; the mean over all of them is what is reported
        e2e_steps = args.steps if ms > 5.0 else max(args.steps, 100)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ta = time.perf_counter()
            e2e_step(mode)
            per_step.append(time.perf_counter() - ta)
        torch.cuda.synchronize()
        e2e_by_mode[mode] = max_over_ranks((time.perf_counter() - t0) * 1e3 / e2e_steps)
        e2e_spread = {"steps": e2e_steps, "median_ms": max_over_ranks(float(np.median(per_step)) * 1e3), "max_ms": max_over_ranks(float(np.max(per_step)) * 1e3)}
        if mode == e2e_modes[0]:
            sampler.mark(False)
        gc.enable()
        e2e_recall_m = recall(kh.numpy().view(np.uint64), truth)
        if mode == e2e_modes[0]:
            e2e_recall = e2e_recall_m
    e2e_mode = e2e_modes[0]                 # the headline is the path the package takes by default, whatever the other one measured
    # where an end-to-end step goes (not timed into anything): host time until everything is enqueued, device time from the first
    # launch to the merged result and to the end of the copies, wall time of the whole step - per rank, max and mean over ranks
    e2e_breakdown = None
    if os.environ.get("VSANN_BENCH_E2E_TRACE", "1" if world > 1 else "0") != "0" and world > 1 and p2p:
        e2e_breakdown = {}
        names = ("host_enqueue", "device_first_launch_to_merged", "device_to_end_of_copies", "wall_step")
        for mode in e2e_modes:
            acc = np.zeros(4)
            n_tr = max(10, args.steps)
            barrier()
            for it in range(n_tr + 3):
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                t0 = time.perf_counter()
                ev[0].record(ts)
                i = p2p.search(g, nq, qh.data_ptr(), ef=chosen, stream=stream)
                p2p.merge(i, mk, md)
                ev[1].record(ts)
                kh.copy_(mk, non_blocking=True); dh.copy_(md, non_blocking=True)
                ev[2].record(ts)
                t1 = time.perf_counter()
                ts.synchronize()
                t2 = time.perf_counter()
                if it >= 3:
                    acc += np.array([(t1 - t0) * 1e3, ev[0].elapsed_time(ev[1]), ev[0].elapsed_time(ev[2]), (t2 - t0) * 1e3])
            acc /= n_tr
            tmax = torch.tensor(acc, device=dev); tsum = tmax.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX); dist.all_reduce(tsum)
            e2e_breakdown[mode] = {"ms_max_over_ranks": dict(zip(names, [round(float(x), 4) for x in tmax.cpu()])),
                                   "ms_mean_over_ranks": dict(zip(names, [round(float(x) / world, 4) for x in tsum.cpu()]))}
        if rank == 0:
            log(f"e2e breakdown: {e2e_breakdown}")
    e2e_ms = e2e_by_mode[e2e_mode]
    if world == 1 and rank == 0 and sampler.need_tail():
        # the sampler was too slow for the millisecond-scale regions above (nvidia-smi fallback): keep the timed
        # step itself running (not timed again) until it has seen the clocks under the same load
        sampler.mark(True)
        t_end = time.perf_counter() + 3.0
        while sampler.need_tail() and time.perf_counter() < t_end:
            search_step(chosen)
            torch.cuda.synchronize()
        sampler.mark(False)
    clocks = sampler.stop() if rank == 0 else None

    # kernel-side work counters at the operating point (every rank: the device counters cost atomics, so outside the timed regions)
    g.set_stats(True); g.reset_stats()
    local_search(chosen); torch.cuda.synchronize()
    sk = g.stats()
    g.set_stats(False)
    k_ev, k_ex = sk["search_dist_evals"] / max(1, sk["search_queries"]), sk["search_expansions"] / max(1, sk["search_queries"])
    st_all = g.stats()

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
        row_bytes = 2 * M * 4
        if host_gen and G == 1:
            ev, ex = algorithmic_counts(vst, g, x, q, chosen, k)
            alg_src = "oracle beam (ovs_graph_search) on the GPU-built graph, 200-query sample"
        else:
            ev, ex = k_ev, k_ex
            alg_src = "the kernel's own counters (every evaluation it made, re-evaluations after visited-hash evictions included): the graph is too large to export to the host oracle"
        alg_bytes = (ev * d * 4 + ex * row_bytes) * nq
        achieved = alg_bytes / (ms_kernel * 1e-3) / 1e9

        # ---- CPU baseline on the same corpus (c2: all of it), the same queries, the same ef; parity at the bench configuration ----
        cb, parity_at_bench = None, None
        build_alg = None
        if not args.no_cpu:
            log("CPU baseline: building the oracle index ...")
            if host_gen and S == 1:
                cpu_n = args.cpu_n or n
                cpu = CpuIndex(vst, d, metric_name, k, cpu_n)
                cpu.add(keys[:cpu_n], x[:cpu_n])
                cq, ctruth = q, truth
                if cpu_n != n:
                    ctruth = numpy_truth(x[:cpu_n], keys[:cpu_n], q[:1000], k, metric_name); cq = q[:1000]
                sample = (f"oracle HNSW (M={M}, ef_add={EF_ADD}, AVX2 distances={cpu.simd}) built on {cpu_n} of the {n} vectors of the workload with "
                          f"{cpu.cores} threads; every step = the workload's {nq} queries, k={k}")
            else:
                cpu_n = args.cpu_n or (min(n, 1_000_000) if host_gen else 200_000)
                if host_gen:
                    xs = x[:cpu_n]
                else:
                    # the first cpu_n rows of THIS shard: regenerate its first chunk with the same seed and chunk size (the device generator's
                    # values depend on the size of the draw), so that corpus and queries come from the same distribution and cluster centres
                    first = DeviceRows(torch, dev, d, 42 + 1000 * rank, args.dist, metric_name in ("COS", "IP")).next(min(n_local, 4_000_000))
                    xs = first[:cpu_n].cpu().numpy().copy()
                    del first
                    torch.cuda.empty_cache()
                cpu = CpuIndex(vst, d, metric_name, k, cpu_n)
                ckeys = np.arange(cpu_n, dtype=np.uint64)
                cpu.add(ckeys, xs)
                cq = q[:1000]
                ctruth = numpy_truth(xs, ckeys, cq, k, metric_name)
                sample = (f"oracle HNSW (M={M}, ef_add={EF_ADD}, AVX2 distances={cpu.simd}) on a {cpu_n}-vector SUBSAMPLE (the first rows of rank 0's shard) with "
                          f"{cpu.cores} threads; every step = {nq} queries, k={k}")
            log(f"CPU baseline: built {cpu.n} vectors in {cpu.build_s:.1f} s")
            ctab, cpick = cpu.sweep(cq, ctruth, grid)
            c_at_gpu_ef = recall(cpu.o.search(cq, k, ef=chosen)[0], ctruth)
            t_same = cpu.time_steps(q, chosen, 3, 1, budget_s=30.0)
            t_own = cpu.time_steps(q, cpick, 3, 1, budget_s=30.0) if (cpick and cpick != chosen) else t_same
            cev, cex = cpu.work(cq[:1000], chosen)
            cst = cpu.o.stats()
            cb = {"value": nq / float(np.mean(t_same)), "unit": "queries/s", "cores": cpu.cores, "kind": "port", "sample": sample,
                  "ef_search": chosen, "recall_at_k": round(c_at_gpu_ef, 4), "ms_per_step": 1e3 * float(np.mean(t_same)), "steps": len(t_same),
                  "at_own_smallest_ef": {"ef_search": cpick, "recall_at_k": ctab.get(cpick), "value": nq / float(np.mean(t_own))},
                  "recall_by_ef": ctab, "build_vectors_per_s": cpu.n / cpu.build_s, "build_seconds": cpu.build_s,
                  "build_dist_evals_per_vector": cst["build_dist_evals"] / max(1, cpu.n), "dist_evals_per_query": cev, "expansions_per_query": cex}
            if host_gen and S == 1 and cpu.n == n:
                g_at = sweep.get(chosen)
                parity_at_bench = {"ef_search": chosen, "gpu_recall_at_k": g_at, "oracle_recall_at_k": round(c_at_gpu_ef, 4),
                                   "gpu_minus_oracle": round(g_at - c_at_gpu_ef, 4), "pass": bool(g_at >= c_at_gpu_ef - 0.01),
                                   "at_parity_point": {"ef_search": parity_ef, "gpu_recall_at_k": sweep.get(parity_ef),
                                                       "oracle_recall_at_k": round(recall(cpu.o.search(cq, k, ef=parity_ef)[0], ctruth), 4)},
                                   "definition": f"same {n} vectors, same {nq} queries, same M/ef_add/ef, exact ground truth from the GPU's bit-exact brute force; pass = gpu >= oracle - 0.01"}
            build_alg = cst["build_dist_evals"] / max(1, cpu.n)
            cpu.close()
        if build_alg is None:
            build_alg = st_all["build_dist_evals"] / max(1, st_all["build_vectors"]) if st_all.get("build_dist_evals") else None
            build_alg_src = "kernel counters" if build_alg else None
        else:
            build_alg_src = "oracle (sequential HNSW insert, same data and parameters): build_dist_evals per vector"
        if build_alg:
            # algorithmic bytes per inserted vector: every distance evaluation of the sequential algorithm gathers one row (d x 4 B);
            # link rows: one level-0 row written + M reverse rows read and rewritten (2M x 4 B each)
            bpv = build_alg * d * 4 + (1 + 2 * M) * row_bytes
            b_ach = bpv * n / build["seconds_wall"] / 1e9      # per GPU
            build["roofline"] = {"bound": "hbm", "kernels": "k_build_search_t + k_select + k_reverse_merge + batch-kNN", "achieved": b_ach, "peak": hbm_peak,
                                 "unit": "GB/s", "frac": b_ach / hbm_peak, "bytes_per_vector": bpv, "dist_evals_per_vector": build_alg,
                                 "source": build_alg_src, "denominator": "wall seconds of the build (host rows, H2D inside)", "peak_source": peak_src}

        traffic = None
        try:   # DRAM bytes per launch of the search kernel from the newest committed ncu --set full capture at this ef (profiles/)
            tj = json.load(open(os.path.join(ROOT, "profiles", "r02_k_search_traffic.json")))
            if S == 1 and tj.get("ef") == chosen and tj.get("config", "c2") == args.config and tj.get("dist", "clustered") == args.dist:
                traffic = tj["dram_bytes_per_launch"]
        except Exception:
            pass
        tf32_peak = peaks.get("tf32_tflops") or peaks.get("bf16_tflops", 1590.0) / 2.0
        if exact is None and not args.no_extra:
            scans = max(1, st_exact["exact_tc_scans"])
            scan_ms = 1e3 * st_exact["exact_tc_scan_seconds"] / scans
            exact_tf = 2.0 * nq * n * d / (scan_ms * 1e-3) / 1e12 if scan_ms > 0 else None
            exact = {"mode": "brute force, bit-identical ids and distances (tcgen05 TF32 scan + exact FP32 re-rank + proof)",
                     "qps": nq / (ms_exact * 1e-3), "ms_per_batch": ms_exact,
                     "tc_queries": st_exact["exact_tc_queries"], "second_scan_queries": st_exact["exact_tc_second_scan"],
                     "fp32_fallback_queries": st_exact["exact_tc_fallback"], "by_batch": exact_by_batch,
                     "roofline": {"bound": "tensor", "kernel": "k_tc_scan<EMIT>", "achieved": exact_tf, "peak": tf32_peak, "unit": "TFLOP/s",
                                  "frac": (exact_tf / tf32_peak) if exact_tf else None, "kernel_ms": scan_ms,
                                  "peak_source": "TF32 = half the measured bf16 burst rate (MEASURED_PEAKS.json bf16_tflops / 2)",
                                  "tf32_tflops_measured_here": tf32_measured,
                                  "flops_per_launch": 2.0 * nq * n * d}}
        extra = {}
        if host_gen and S == 1 and not args.no_extra and args.dist == "clustered":
            extra["iid_gaussian"] = iid_block(args, vst, torch, lib, local, dev, ts, cfg, not args.no_cpu)
        if S == 1 and not args.no_extra and not args.no_cpu and args.config == "c2":
            extra["via_actor"] = via_actor_leg(vst, True)
        total = n * S
        line = {
            "metric": "ann_qps_at_recall95", "value": nq / (ms * 1e-3), "unit": "queries/s", "n_gpus": S, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_dict(args, cfg, S),
            "process_model": "one process, one vs_index over %d devices" % G if G > 1 else ("one process per GPU (torchrun)" if world > 1 else "one process, one GPU"),
            "operating_point": {"ef_search": chosen, "recall_at_k": sweep.get(chosen), "recall_target_reached": reached,
                                "criterion": "smallest ef reaching the target on a shard searched alone, max over ranks (the same criterion at every N)",
                                "ef_sweep_recall_shard_alone": sweep_local, "ef_sweep_recall": sweep,
                                "smallest_ef_global": {"ef_search": ef_global, "recall_at_k": sweep.get(ef_global), "ms_per_step": ms_global_ef,
                                                       "qps": nq / (ms_global_ef * 1e-3)},
                                "exchange": exchange,
                                "timed_step_start": ("ranks aligned by a device-side barrier kernel before the first event (VSANN_BENCH_ALIGN=0: NCCL barrier only)"
                                                     if (world > 1 and p2p is not None and align_starts) else "barrier + synchronize")},
            "value_definition": ("queries answered per second against the whole corpus of corpus_total vectors "
                                 "(weak scaling: the corpus grows N x at fixed per-GPU work, so the ideal curve is FLAT; see shard_queries_per_s)"
                                 if cfg["scaling"] == "weak" else "queries answered per second against the whole corpus (split over the ranks)"),
            "shard_queries_per_s": S * nq / (ms * 1e-3), "vectors_searched_per_s": total * nq / (ms * 1e-3),
            "parity_point": {"ef_search": parity_ef, "recall_at_k": sweep.get(parity_ef), "qps": nq / (ms_parity * 1e-3)},
            "parity_at_bench": parity_at_bench,
            "build": build,
            "e2e": {"value": nq / (e2e_ms * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": nq * d * 4 * world, "d2h_bytes_per_step": nq * k * 12 * world,
                    "ms_per_step": e2e_ms, "recall_at_k": e2e_recall, "path": e2e_mode, "per_step_spread": e2e_spread, "breakdown": e2e_breakdown},
            "gpu_launches": args.steps * (1 if S == 1 else (G + 1 if G > 1 else 2)),   # per step: k_search_t per shard (+ the merge kernel after the exchange)
            "by_batch": by_batch or None, "recall_sweep": recall_sweep,
            "exact": exact,
            "roofline": {"bound": "hbm", "kernel": "k_search_t", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel_ms": ms_kernel,
                         "algorithmic": {"dist_evals_per_query": ev, "expansions_per_query": ex, "bytes_per_query": ev * d * 4 + ex * row_bytes,
                                         "source": alg_src},
                         "kernel_counted": {"dist_evals_per_query": k_ev, "expansions_per_query": k_ex}},
            "cpu_baseline": cb, "clocks": clocks,
        }
        line.update(extra)
        emit(line)
    g.close()
    if world > 1:
        barrier()
        dist.destroy_process_group()


def iid_block(args, vst, torch, lib, local, dev, ts, cfg, with_cpu):
    """SURVEY.md §8d's primary generator for configs[1] - i.i.d. standard_normal, adversarial for any graph index (intrinsic
    dimension 128) - beside the clustered default: same sizes, same parameters, its own ef sweep, GPU and oracle side by side."""
    d, k, n, nq = cfg["d"], cfg["k"], args.n or cfg["n"], args.nq or cfg["nq"]
    stream = ts.cuda_stream
    x = gen(n, d, 42, "gaussian"); q = gen(nq, d, 43, "gaussian")
    keys = np.arange(n, dtype=np.uint64)
    g = vst.VsIndex(d, metric=vst.Metric.L2SQ, connectivity=M, expansion_add=EF_ADD, expansion_search=EF_PARITY, capacity_hint=n, device=local, lib=lib)
    xh = torch.from_numpy(x).pin_memory()
    t0 = time.perf_counter()
    for a in range(0, n, 1 << 18):
        g.add_batch(keys[a:a + (1 << 18)], xh[a:a + (1 << 18)].numpy())
    assert g.count() == n
    bw = time.perf_counter() - t0
    del xh
    qd = torch.from_numpy(q).to(dev)
    ok = torch.empty((nq, k), dtype=torch.int64, device=dev); od = torch.empty((nq, k), dtype=torch.float32, device=dev)

    def run(ef, exact=False):
        g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), ef=ef, exact=exact, stream=stream)
        torch.cuda.synchronize()
        return ok.cpu().numpy().view(np.uint64).copy()

    truth = run(0, exact=True)
    sweep, chosen = {}, None
    for ef in EF_GRID[k]:
        r = recall(run(ef), truth)
        sweep[ef] = round(r, 4)
        if r >= RECALL_TARGET:
            chosen = ef
            break
    if EF_PARITY not in sweep:
        sweep[EF_PARITY] = round(recall(run(EF_PARITY), truth), 4)
    efq = chosen or max(sweep)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot, steps = 0.0, max(3, args.steps // 2)
    run(efq)
    for _ in range(steps):
        flush_buf.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ts)
        g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), ef=efq, stream=stream)
        e1.record(ts); e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / steps
    out = {"workload": f"{n} x {d} f32 L2sq i.i.d. standard_normal (numpy PCG64 seeds 42/43), M={M} ef_add={EF_ADD}, {nq} queries k={k}",
           "build_vectors_per_s": n / bw, "build_seconds_wall": bw, "ef_sweep_recall": sweep, "ef_search": efq, "recall_target_reached": chosen is not None,
           "recall_at_k": sweep.get(efq), "qps": nq / (ms * 1e-3), "ms_per_step": ms}
    g.close()
    del qd
    if with_cpu:
        cpu = CpuIndex(vst, d, "L2SQ", k, n)
        cpu.add(keys, x)
        efs = sorted({EF_PARITY, efq})
        out["oracle"] = {"build_vectors_per_s": cpu.n / cpu.build_s, "cores": cpu.cores,
                         "recall_by_ef": {str(e): round(recall(cpu.o.search(q, k, ef=e)[0], truth), 4) for e in efs}}
        t = cpu.time_steps(q, efq, 2, 1, budget_s=30.0)
        out["oracle"]["qps"] = nq / float(np.mean(t)); out["oracle"]["ef_search"] = efq
        go, oo = sweep.get(efq), out["oracle"]["recall_by_ef"][str(efq)]
        out["parity"] = {"ef_search": efq, "gpu_recall_at_k": go, "oracle_recall_at_k": oo, "gpu_minus_oracle": round(go - oo, 4), "pass": bool(go >= oo - 0.01)}
        cpu.close()
    return out


if __name__ == "__main__":
    main()
