"""GPU parity of the streaming exact scan (csrc/exact_kernels.cu k_exact_stream: batches of <= 16 queries, one
thread per corpus row, TMA-staged tiles, sequential fmaf chains) against the CPU oracle: ids AND distance bits
identical for every metric, dimension, batch size and k; ties, tombstones and replaces included. The stats prove
which path answered."""
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


@pytest.mark.parametrize("metric,n,d,k", [(1, 40000, 128, 10), (0, 30011, 128, 10), (2, 25000, 96, 100), (0, 9000, 100, 128),
                                          (0, 6000, 768, 10), (2, 5000, 1536, 100), (1, 66000, 20, 1), (1, 3000, 2048, 200),
                                          (2, 300, 64, 10), (1, 257, 36, 256)])
def test_stream_exact_bit_identical(vst, oracle_lib, monkeypatch, metric, n, d, k):
    monkeypatch.setenv("VSANN_EXACT_STREAM", "1")
    x = make_data(n, d, 200 + n); q = make_data(16, d, 201 + n)
    g, o = pair(vst, oracle_lib, d, metric, n)
    keys = np.random.default_rng(n).permutation(n).astype(np.uint64) * 5 + 11     # slot order != key order
    g.add_batch(keys, x); o.add_batch(keys, x)
    total = 0
    for nq in (1, 2, 3, 5, 8, 13, 16):          # every pass shape: 1, 2 and 4 query pairs, ragged last pass
        same(g, o, q[:nq], k)
        total += nq
    st = g.stats()
    assert st["exact_stream_queries"] == total, "the streaming scan must have answered"
    assert st["exact_tc_queries"] == 0
    g.close()


def test_stream_exact_ties_tombstones_replaces(vst, oracle_lib, monkeypatch):
    """Exact duplicates (ties broken by key), 1-ulp neighbours, removed and replaced keys, queries on top of rows."""
    monkeypatch.setenv("VSANN_EXACT_STREAM", "1")
    n, d = 36000, 64
    x = make_data(n, d, 31)
    x[6000:12000] = x[0:6000]                                   # duplicates
    x[12000:18000] = np.nextafter(x[0:6000], np.float32(np.inf))  # one ulp away
    q = np.concatenate([x[:8], make_data(8, d, 32)])
    g, o = pair(vst, oracle_lib, d, 1, 2 * n)
    keys = np.random.default_rng(7).permutation(n).astype(np.uint64)
    for ix in (g, o):
        ix.add_batch(keys, x)
        for kk in keys[:300]:
            assert ix.remove(int(kk))
        ix.add_batch(keys[20000:20300], make_data(300, d, 33))   # replaces (old slots become tombstones)
    assert g.count() == o.count() == n - 300
    same(g, o, q, 25)
    same(g, o, q[:1], 40)
    assert g.stats()["exact_stream_queries"] == 17
    g.close()


def test_stream_massive_ties_and_short_corpus(vst, oracle_lib, monkeypatch):
    """Thousands of equal rows (every threshold comparison is a tie) and fewer live rows than k."""
    monkeypatch.setenv("VSANN_EXACT_STREAM", "1")
    n, d = 20000, 32
    x = np.tile(make_data(4, d, 41), (n // 4, 1))               # 4 distinct vectors, 5000 copies each
    q = np.concatenate([x[:4], make_data(4, d, 42)])
    g, o = pair(vst, oracle_lib, d, 0, n)
    keys = np.arange(n, dtype=np.uint64)[::-1].copy()
    g.add_batch(keys, x); o.add_batch(keys, x)
    same(g, o, q, 10)
    g.close()
    g, o = pair(vst, oracle_lib, 24, 1, 1000)
    x = make_data(7, 24, 43)
    g.add_batch(np.arange(7, dtype=np.uint64), x); o.add_batch(np.arange(7, dtype=np.uint64), x)
    same(g, o, make_data(3, 24, 44), 10)                        # found == 7 < k
    g.close()


def test_stream_selected_by_size(vst, oracle_lib, monkeypatch):
    """Default policy: <= 16 queries against a corpus of >= 2^22 floats stream; 17 queries do not."""
    monkeypatch.delenv("VSANN_EXACT_STREAM", raising=False)
    n, d = 40000, 128
    x = make_data(n, d, 51); q = make_data(17, d, 52)
    g, o = pair(vst, oracle_lib, d, 1, n)
    keys = np.arange(n, dtype=np.uint64)
    g.add_batch(keys, x); o.add_batch(keys, x)
    same(g, o, q[:16], 10)
    assert g.stats()["exact_stream_queries"] == 16
    same(g, o, q, 10)
    assert g.stats()["exact_stream_queries"] == 16
    monkeypatch.setenv("VSANN_EXACT_STREAM", "0")
    same(g, o, q[:4], 10)
    assert g.stats()["exact_stream_queries"] == 16
    g.close()
