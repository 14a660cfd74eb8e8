"""GPU-vs-oracle recall probe (dev tool): python scripts/recall_probe.py [n] [d]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import vector_store_text_b200 as vst
from conftest import make_data, recall_at_k
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
olib = vst.load_library(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "liboracle_vsann.so"), prefix="ovs_")
for metric, clustered in [(1, False), (1, True), (0, True), (0, False), (2, True)]:
    x = make_data(n, d, 31, clustered=clustered); q = make_data(500, d, 32, clustered=clustered)
    if metric == 2:
        x /= np.linalg.norm(x, axis=1, keepdims=True); q /= np.linalg.norm(q, axis=1, keepdims=True)
    keys = np.arange(n, dtype=np.uint64)
    o = vst.VsIndex(d, metric=metric, capacity_hint=n, lib=olib); o.set_threads(os.cpu_count()); o.add_batch(keys, x)
    truth, _, _ = o.search(q, 10, exact=True)
    ro = {ef: recall_at_k(o.search(q, 10, ef=ef)[0], truth) for ef in (32, 64, 128)}
    line = f"metric={metric} clustered={clustered} oracle={ro}"
    for mb in (2, 4, 8, 16):
        os.environ['VSANN_BATCH_DIV'] = str(mb)
        g = vst.VsIndex(d, metric=metric, capacity_hint=n)
        t0 = time.time(); g.add_batch(keys, x); g.flush(); tb = time.time() - t0
        rg = {ef: round(recall_at_k(g.search(q, 10, ef=ef)[0], truth), 4) for ef in (32, 64, 128)}
        links, _ = g.export_level(0)
        deg = (links[:n] != 0xFFFFFFFF).sum(1).mean()
        line += f"\n   gpu batch_div={mb}: {rg} build={tb:.2f}s deg0={deg:.1f}"
        g.close()
    print(line, flush=True)
