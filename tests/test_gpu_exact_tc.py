"""GPU parity of the tensor-core exact search (csrc/exact_tc.cu: tcgen05 TF32 scan + exact FP32 re-rank with
proof) against the CPU oracle: ids AND distance bits identical, whatever the metric, dimension, k, ties,
tombstones or replaces; the stats prove which path answered."""
import numpy as np
import pytest

from conftest import make_data

pytestmark = pytest.mark.gpu


def pair(vst, oracle_lib, d, metric, cap):
    kw = dict(connectivity=2, expansion_add=8)      # exact search ignores the graph: keep both builds cheap
    g = vst.VsIndex(d, metric=metric, capacity_hint=cap, **kw)
    o = vst.VsIndex(d, metric=metric, capacity_hint=cap, lib=oracle_lib, **kw)
    o.set_threads(16)
    return g, o


def same(g, o, q, k):
    gk, gd, gf = g.search(q, k, exact=True)
    ok, od, of = o.search(q, k, exact=True)
    assert np.array_equal(gf, of)
    assert np.array_equal(gk, ok), "ids must be bit-identical to the oracle's exact search"
    assert np.array_equal(gd.view(np.uint32), od.view(np.uint32)), "distances must carry the oracle's bits"


@pytest.mark.parametrize("metric,n,d,k,nq", [(1, 40000, 128, 10, 300), (0, 40000, 128, 10, 129), (2, 50000, 96, 100, 200),
                                             (0, 33000, 100, 128, 150), (0, 20000, 768, 10, 64), (2, 17000, 1536, 100, 40),
                                             (1, 66000, 20, 1, 500)])
def test_tc_exact_bit_identical(vst, oracle_lib, monkeypatch, metric, n, d, k, nq):
    monkeypatch.setenv("VSANN_EXACT_TC", "1")
    x = make_data(n, d, 100 + n); q = make_data(nq, d, 101 + n)
    g, o = pair(vst, oracle_lib, d, metric, n)
    keys = np.random.default_rng(n).permutation(n).astype(np.uint64) * 7 + 3     # slot order != key order
    g.add_batch(keys, x); o.add_batch(keys, x)
    same(g, o, q, k)
    st = g.stats()
    assert st["exact_tc_queries"] == nq, "the tensor-core path must have answered"
    assert st["exact_tc_fallback"] == 0
    g.close()


def test_tc_exact_ties_tombstones_replaces(vst, oracle_lib, monkeypatch):
    """Exact duplicates (ties broken by key), 1-ulp neighbours, removed and replaced keys, queries on top of rows."""
    monkeypatch.setenv("VSANN_EXACT_TC", "1")
    n, d = 36000, 64
    x = make_data(n, d, 31)
    x[6000:12000] = x[0:6000]                                   # duplicates
    x[12000:18000] = np.nextafter(x[0:6000], np.float32(np.inf))  # one ulp away
    q = np.concatenate([x[:150], make_data(150, d, 32)])
    g, o = pair(vst, oracle_lib, d, 1, 2 * n)
    keys = np.random.default_rng(7).permutation(n).astype(np.uint64)
    for ix in (g, o):
        ix.add_batch(keys, x)
        for kk in keys[100:400]:
            assert ix.remove(int(kk))
        ix.add_batch(keys[20000:20300], make_data(300, d, 33))   # replaces (old slots become tombstones)
    assert g.count() == o.count() == n - 300
    same(g, o, q, 25)
    assert g.stats()["exact_tc_queries"] == 300
    g.close()


def test_tc_exact_massive_ties_fall_back_correctly(vst, oracle_lib, monkeypatch):
    """More equal rows than the candidate buffer holds: the proof cannot succeed, the FP32 kernel answers."""
    monkeypatch.setenv("VSANN_EXACT_TC", "1")
    n, d, k = 20000, 32, 10
    x = np.tile(make_data(4, d, 41), (n // 4, 1))               # 4 distinct vectors, 5000 copies each
    q = np.concatenate([x[:8], make_data(24, d, 42)])
    g, o = pair(vst, oracle_lib, d, 0, n)
    keys = np.arange(n, dtype=np.uint64)
    g.add_batch(keys, x); o.add_batch(keys, x)
    same(g, o, q, k)
    st = g.stats()
    assert st["exact_tc_queries"] == 32 and st["exact_tc_fallback"] > 0
    g.close()


def test_tc_selected_by_size_and_skipped_for_large_k(vst, oracle_lib, monkeypatch):
    monkeypatch.delenv("VSANN_EXACT_TC", raising=False)
    n, d = 70000, 32
    x = make_data(n, d, 51); q = make_data(256, d, 52)
    g, o = pair(vst, oracle_lib, d, 1, n)
    keys = np.arange(n, dtype=np.uint64)
    g.add_batch(keys, x); o.add_batch(keys, x)
    same(g, o, q, 10)                                           # 70000 x 256 pairs >= 2^24: tensor path
    assert g.stats()["exact_tc_queries"] == 256
    same(g, o, q[:16], 200)                                     # k > 128: FP32 tile kernel
    same(g, o, q[:8], 10)                                       # tiny batch: FP32 tile kernel
    assert g.stats()["exact_tc_queries"] == 256
    monkeypatch.setenv("VSANN_EXACT_TC", "0")
    same(g, o, q, 10)
    assert g.stats()["exact_tc_queries"] == 256
    g.close()


def test_exact_search_on_a_user_stream_does_not_synchronise(vst, oracle_lib, monkeypatch):
    """include/vsann.h: vs_search_batch_device(stream != NULL) returns without synchronising - in exact mode too: the second scan and
    the fallback are driven by counters that stay on the device. The call must come back while the GPU is still working."""
    import torch
    monkeypatch.setenv("VSANN_EXACT_TC", "1")
    n, d, nq, k = 400000, 128, 8192, 10
    x = make_data(n, d, 61); q = make_data(nq, d, 62)
    g = vst.VsIndex(d, metric=1, capacity_hint=n, connectivity=2, expansion_add=8)
    g.add_batch(np.arange(n, dtype=np.uint64), x)
    dev = torch.device("cuda", 0)
    qd = torch.from_numpy(q).to(dev)
    ok = torch.empty((nq, k), dtype=torch.int64, device=dev); od = torch.empty((nq, k), dtype=torch.float32, device=dev)
    st = torch.cuda.Stream(device=dev)
    g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), exact=True, stream=st.cuda_stream)   # warm-up (scratch allocation)
    st.synchronize()
    pending = 0
    for _ in range(5):
        done = torch.cuda.Event()
        g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), exact=True, stream=st.cuda_stream)
        done.record(st)
        pending += not done.query()          # still running when the call has returned
        st.synchronize()
    assert pending >= 4, "the exact path must not wait for the GPU"
    want, _, _ = g.search(q[:64], k, exact=True)
    assert np.array_equal(ok.cpu().numpy().view(np.uint64)[:64], want)
    g.close()
