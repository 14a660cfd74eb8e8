"""GPU parity on the BASELINE.json config shapes at reduced N (the full sizes are bench workloads): every
dimension / metric / k combination the configs name goes through build (K4/K5), ANN search (K3: the register
-chunk variants CPL=4 for d=128, CPL=3 for d=96, the generic variant with 3 KB / 6 KB bulk-copied rows for
d=768 / 1536, list sizes NL=1..8 and the shared-memory list above ef=256) and exact search, against the oracle."""
import numpy as np
import pytest

from conftest import make_data, recall_at_k

pytestmark = pytest.mark.gpu

SHAPES = [
    # name, n, d, metric, k, nq, efs
    ("c1-10k-128-cos", 10000, 128, 0, 10, 1000, [64]),
    ("c2-1M-128-l2", 30000, 128, 1, 10, 500, [24, 64, 128]),
    ("c3-10M-768-cos", 8000, 768, 0, 10, 256, [64]),
    ("c4-100M-96-l2", 20000, 96, 1, 10, 500, [32, 64]),
    ("c5-5M-1536-ip", 6000, 1536, 2, 100, 128, [100, 128, 256, 512]),
]


@pytest.mark.parametrize("name,n,d,metric,k,nq,efs", SHAPES, ids=[s[0] for s in SHAPES])
def test_config_shape(vst, oracle_lib, name, n, d, metric, k, nq, efs):
    x = make_data(n, d, 61 + d, clustered=True, centres=256); q = make_data(nq, d, 62 + d, clustered=True, centres=256)
    if metric != 1:   # rows L2-normalised for the cosine / dot-product configs
        x /= np.linalg.norm(x, axis=1, keepdims=True); q /= np.linalg.norm(q, axis=1, keepdims=True)
    kw = dict(connectivity=16, expansion_add=128, expansion_search=64)
    g = vst.VsIndex(d, metric=metric, capacity_hint=n, **kw)
    o = vst.VsIndex(d, metric=metric, capacity_hint=n, lib=oracle_lib, **kw)
    o.set_threads(16)
    keys = np.arange(n, dtype=np.uint64)
    g.add_batch(keys, x); o.add_batch(keys, x)
    assert g.count() == o.count() == n
    # brute force: bit-identical
    gk, gd, _ = g.search(q, k, exact=True)
    ok, od, _ = o.search(q, k, exact=True)
    assert np.array_equal(gk, ok) and np.array_equal(gd.view(np.uint32), od.view(np.uint32))
    # ANN: recall not below the oracle's HNSW at the same M / ef_add / ef
    for ef in efs:
        ak, ad, _ = g.search(q, k, ef=ef)
        bk, _, _ = o.search(q, k, ef=ef)
        rg, ro = recall_at_k(ak, gk), recall_at_k(bk, gk)
        print(f"{name} ef={ef} recall@{k} gpu={rg:.4f} oracle={ro:.4f}")
        assert rg >= ro - 0.02, (name, ef, rg, ro)
        assert np.all(np.diff(ad, axis=1) >= 0)
        assert np.all(ak != np.uint64(0xFFFFFFFFFFFFFFFF))
    g.close()


def test_c1_via_the_native_index_actor(vst):
    """BASELINE configs[0]: 10k x 128-d cosine, add + 1k ANN queries k=10 *via the index actor* — here the C++ actor
    (csrc/actor.cpp: channel 10, 2 x T in flight, bimap) over the GPU library, one message per vector / query."""
    import time
    n, d, nq, k = 10000, 128, 1000, 10
    x = make_data(n, d, 71, clustered=True, centres=256); q = make_data(nq, d, 72, clustered=True, centres=256)
    a = vst.NativeActor("vector.ann", d, metric=vst.Metric.COS, capacity_hint=n, threads=8)
    for i in range(n):
        a.add_or_replace(("pk", i), x[i])
    t0 = time.time()
    while a.count() != n and time.time() - t0 < 60:
        time.sleep(0.01)
    assert a.count() == n
    truth_ix = vst.VsIndex(d, metric=vst.Metric.COS, capacity_hint=n, connectivity=2, expansion_add=8)
    truth_ix.add_batch(np.arange(n, dtype=np.uint64), x)
    truth, _, _ = truth_ix.search(q, k, exact=True)
    hits = 0
    for i in range(nq):
        pks, dist = a.ann(q[i], k)
        assert len(pks) == k and all(dist[j] <= dist[j + 1] for j in range(k - 1))
        hits += len({p[1] for p in pks} & set(truth[i].tolist()))
    rec = hits / (nq * k)
    print(f"c1 via native actor: recall@10 = {rec:.4f}")
    assert rec >= 0.80
    st = a.stats()
    assert st["keys"] == n and st["max_queued"] <= 10 and st["max_in_flight"] <= 16
    a.close(); truth_ix.close()
