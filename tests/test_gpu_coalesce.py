"""vs_search from many threads at once (the reference runs up to 2 x T searches concurrently, usearch.rs:118-131,275-276):
concurrent single-query calls share launches, every caller still gets exactly the answer a lone call gets."""
import threading

import numpy as np
import pytest

from conftest import make_data

pytestmark = pytest.mark.gpu


def hammer(ix, q, ks, threads):
    out = [None] * len(q)
    err = []

    def work(t):
        try:
            for i in range(t, len(q), threads):
                out[i] = ix.search_one(q[i], ks[i])
        except Exception as e:   # noqa: BLE001
            err.append(e)

    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not err, err
    return out


def test_concurrent_vs_search_calls_share_launches(vst):
    n, d, nq = 20000, 64, 1024
    x = make_data(n, d, 11, clustered=True); q = make_data(nq, d, 12, clustered=True)
    g = vst.VsIndex(d, metric=vst.Metric.L2SQ, capacity_hint=n)
    g.add_batch(np.arange(n, dtype=np.uint64), x)
    ks = [10 if i % 3 else 5 for i in range(nq)]              # two different k in flight at once
    want10, wd10, _ = g.search(q, 10)
    want5, wd5, _ = g.search(q, 5)
    g.reset_stats()
    got = hammer(g, q, ks, threads=16)
    for i in range(nq):
        wk, wd = (want10, wd10) if ks[i] == 10 else (want5, wd5)
        assert np.array_equal(got[i][0], wk[i]) and np.array_equal(got[i][1], wd[i])
    st = g.stats()
    assert st["search_host_queries"] == nq
    assert st["search_host_launches"] < nq, "16 concurrent callers must have shared launches"
    print(f"coalescer: {nq} calls in {st['search_host_launches']} launches")
    # one caller alone: a launch per call, same answers
    g.reset_stats()
    for i in range(8):
        kk, dd = g.search_one(q[i], 10)
        assert np.array_equal(kk, want10[i]) and np.array_equal(dd, wd10[i])
    st = g.stats()
    assert st["search_host_queries"] == 8 and st["search_host_launches"] == 8
    with pytest.raises(vst.VsError):
        g.search_one(q[0][:10], 10)
    g.close()


def test_coalescer_can_be_switched_off(vst, monkeypatch):
    monkeypatch.setenv("VSANN_COALESCE", "0")
    n, d, nq = 5000, 32, 256
    x = make_data(n, d, 21); q = make_data(nq, d, 22)
    g = vst.VsIndex(d, metric=vst.Metric.COS, capacity_hint=n)
    g.add_batch(np.arange(n, dtype=np.uint64), x)
    want, wd, _ = g.search(q, 10)
    g.reset_stats()
    got = hammer(g, q, [10] * nq, threads=8)
    for i in range(nq):
        assert np.array_equal(got[i][0], want[i]) and np.array_equal(got[i][1], wd[i])
    st = g.stats()
    assert st["search_host_queries"] == nq and st["search_host_launches"] == nq
    g.close()


def test_concurrent_ann_through_the_native_actor(vst):
    """Ann messages from 16 client threads through the C++ actor (channel 10, 2 x T in flight): right answers, shared launches."""
    n, d, nq, k = 8000, 48, 512, 10
    x = make_data(n, d, 31, clustered=True); q = make_data(nq, d, 32, clustered=True)
    a = vst.NativeActor("vector.ann", d, metric=vst.Metric.L2SQ, capacity_hint=n, threads=8)
    for i in range(n):
        a.add_or_replace(("pk", i), x[i])
    import time
    t0 = time.time()
    while a.count() != n and time.time() - t0 < 60:
        time.sleep(0.01)
    assert a.count() == n
    ref = [a.ann(q[i], k) for i in range(nq)]                  # one client: the lone-call answers
    out = [None] * nq

    def work(t):
        for i in range(t, nq, 16):
            out[i] = a.ann(q[i], k)

    th = [threading.Thread(target=work, args=(t,)) for t in range(16)]
    [t.start() for t in th]
    [t.join() for t in th]
    for i in range(nq):
        assert out[i][0] == ref[i][0] and out[i][1] == ref[i][1]
    a.close()
