"""Distribution of per-query work (oracle beam on the GPU-built graph) for two batch schedules: is the slowest query slow because it expands more?"""
import os, sys, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vector_store_text_b200 as vst
n, d, M, ef, NQ = 1_000_000, 128, 16, 24, int(os.environ.get("NQ", 1500))
x = bench.gen(n, d, 42, "clustered"); q = bench.gen(10000, d, 43, "clustered")
olib = bench.load_oracle(vst)
for mb in (32768,):
    os.environ["VSANN_MAX_BATCH"] = str(mb)
    g = vst.VsIndex(d, metric=vst.Metric.L2SQ, connectivity=M, expansion_add=128, expansion_search=64, capacity_hint=n)
    g.add_batch(np.arange(n, dtype=np.uint64), x)
    links0, _ = g.export_level(0)
    deg = (links0 != 0xFFFFFFFF).sum(1)
    uppers = []; lvl = 1
    while M ** lvl < n:
        uppers.append(g.export_level(lvl)[0]); lvl += 1
    evs, exs, dent, dres = [], [], [], []
    slots = np.zeros(10, np.uint32); dist = np.zeros(10, np.float32)
    for i in range(NQ):
        qi = q[i]; cur, dcur = 0, float(((x[0] - qi) ** 2).sum()); ev, ex = 1, 0
        for l in range(len(uppers), 0, -1):
            tab, stride = uppers[l - 1], M ** l
            while True:
                row = tab[cur // stride]; row = row[row != 0xFFFFFFFF]; ex += 1
                if row.size == 0: break
                dd = ((x[row] - qi) ** 2).sum(1); ev += row.size; j = int(dd.argmin())
                if dd[j] < dcur: cur, dcur = int(row[j]), float(dd[j])
                else: break
        found, e1, e2 = C.c_size_t(), C.c_uint64(), C.c_uint64()
        olib.ovs_graph_search(1, d, n, x.ctypes.data_as(C.c_void_p), links0.ctypes.data_as(C.c_void_p), links0.shape[1], None, cur,
                              qi.ctypes.data_as(C.c_void_p), 10, ef, slots.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p),
                              C.byref(found), C.byref(e1), C.byref(e2))
        evs.append(ev + e1.value); exs.append(ex + e2.value); dent.append(dcur); dres.append(float(dist[0]))
    evs, exs = np.array(evs), np.array(exs)
    pc = lambda a: {p: float(np.percentile(a, p)) for p in (50, 90, 99, 100)}
    print(f"max_batch {mb}: degree mean {deg.mean():.2f} full rows {np.mean(deg == 32):.3f} | evals mean {evs.mean():.1f} {pc(evs)} | expansions mean {exs.mean():.1f} {pc(exs)}", flush=True)
    dent, dres = np.array(dent), np.array(dres)
    print('corr(evals, entry dist)', np.corrcoef(evs, dent)[0,1], 'corr(evals, entry/top1)', np.corrcoef(evs, dent/np.maximum(dres,1e-9))[0,1], 'corr(evals, top1 dist)', np.corrcoef(evs, dres)[0,1], 'corr(exps, top1)', np.corrcoef(exs, dres)[0,1])
    np.savez('gpurun_out/tail.npz', evs=evs, exs=exs, dent=dent, dres=dres)
    g.close()
