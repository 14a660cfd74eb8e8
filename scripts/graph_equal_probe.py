"""Builds the same data twice with an env switch flipped and checks that every link table is identical.
usage: SWITCH=VSANN_SELECT_SMEM python scripts/graph_equal_probe.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vector_store_text_b200 as vst
sw = os.environ.get("SWITCH", "VSANN_SELECT_SMEM")
n = int(os.environ.get("N", 200000))
out = []
for d, metric in ((128, 1), (96, 0), (100, 2)):
    x = bench.gen(n, d, 42, "clustered")
    tabs = []
    for v in ("0", "1"):
        os.environ[sw] = v
        g = vst.VsIndex(d, metric=vst.Metric(metric), capacity_hint=n)
        g.add_batch(np.arange(n, dtype=np.uint64), x)
        st = g.stats()
        tabs.append([g.export_level(l)[0] for l in range(st["top_level"] + 1)])
        print(f"d={d} metric={metric} {sw}={v}: build {st['build_seconds']:.3f} s", flush=True)
        g.close()
    same = all(np.array_equal(a, b) for a, b in zip(*tabs))
    print(f"d={d} metric={metric}: link tables identical = {same}", flush=True)
    out.append(same)
sys.exit(0 if all(out) else 1)
