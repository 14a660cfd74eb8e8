"""The callers either side of the index actor (SURVEY.md §8f N2 + N3) over the GPU library (run with -m gpu on a B200):
the reference's integration KAT spoken over HTTP (tests/integration/usearch.rs:20-135) and an out-of-order ingest stream
through the last-write-wins filter (src/monitor_items.rs:56-80) into the native actor, checked against a plain model."""
import time

import numpy as np
import pytest

from conftest import make_data
from test_service import run_integration_kat

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("native", [False, True])
def test_integration_kat_over_http_gpu(vst, native):
    run_integration_kat(vst, vst.load_library(), native)


def test_out_of_order_ingest_through_lww_into_the_gpu_actor(vst):
    """30 % replaces, 5 % deletes, shuffled delivery: what the index answers at the end is exactly the newest live row per key."""
    from vector_store_text_b200.service import DbEmbedding, ItemsMonitor
    d, nkeys, nmsg = 32, 3000, 6000
    rng = np.random.default_rng(5)
    rows = make_data(nmsg, d, 77)
    pk = np.concatenate([np.arange(nkeys), rng.integers(0, nkeys, nmsg - nkeys)])
    ts = rng.permutation(nmsg) + 1                                     # distinct write times
    is_del = rng.random(nmsg) < 0.05
    order = rng.permutation(nmsg)                                      # delivery order != write order
    # ONE worker thread: the actor (like the reference, src/index/usearch.rs:118-131) gives no ordering guarantee between messages in
    # flight - with T workers an accepted "add, remove, add" of one key may run as "add, add, remove" - so a test that predicts the
    # final state must have them processed in order.
    actor = vst.NativeActor("ks.items", d, metric=vst.Metric.L2SQ, capacity_hint=10000, threads=1, lib=vst.load_library())
    mon = ItemsMonitor(actor)
    newest = {}
    for i in order:
        mon.add(DbEmbedding((int(pk[i]),), None if is_del[i] else rows[i], int(ts[i])))
        if pk[i] not in newest or ts[newest[pk[i]]] < ts[i]:
            newest[int(pk[i])] = int(i)
    live = {k: i for k, i in newest.items() if not is_del[i]}
    t0 = time.time()
    while actor.count() != len(live):                                 # adds / removes are fire-and-forget
        assert time.time() - t0 < 20, (actor.count(), len(live))
        time.sleep(0.01)
    assert mon.dropped > 0
    # every live key answers with itself at distance 0 for its newest row; deleted keys never come back
    some = list(live.items())[:200]
    for k, i in some:
        pks, dist = actor.ann(rows[i], 1)
        assert pks == [(k,)] and dist[0] == 0.0
    dead = [k for k, i in newest.items() if is_del[i]][:50]
    for k in dead:
        i = newest[k]
        j = max((m for m in range(nmsg) if pk[m] == k and not is_del[m]), key=lambda m: ts[m], default=None)
        if j is not None:
            pks, _ = actor.ann(rows[j], 5)
            assert (k,) not in pks
    actor.close()
