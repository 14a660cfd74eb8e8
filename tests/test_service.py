"""CPU tests of the callers either side of the index actor, over the oracle backend:
the reference's LWW unit test (src/monitor_items.rs:82-208), the engine registry rules (src/engine.rs:86-123)
and its integration KAT spoken over HTTP (tests/integration/usearch.rs:20-135, tests/integration/httpclient.rs)."""
import json
import time
import urllib.error
import urllib.request

import pytest


class RecordingIndex:
    """The reference fakes the index actor with a bare channel receiver (monitor_items.rs:91,160-206)."""

    def __init__(self):
        self.messages = []

    def add_or_replace(self, pk, emb):
        self.messages.append(("AddOrReplace", pk, list(emb)))

    def remove(self, pk):
        self.messages.append(("Remove", pk))


def test_lww_flow_kat(vst):
    from vector_store_text_b200.service import DbEmbedding, ItemsMonitor
    idx = RecordingIndex()
    mon = ItemsMonitor(idx)
    mon.feed([
        DbEmbedding((1,), [1.], 10),
        DbEmbedding((2,), [2.], 11),
        DbEmbedding((1,), [3.], 5),      # should be dropped
        DbEmbedding((2,), [4.], 15),     # should be accepted
        DbEmbedding((1,), None, 25),
        DbEmbedding((1,), [5.], 24),     # should be dropped
        DbEmbedding((1,), [6.], 26),
    ])
    assert idx.messages == [
        ("AddOrReplace", (1,), [1.]),
        ("AddOrReplace", (2,), [2.]),
        ("AddOrReplace", (2,), [4.]),
        ("Remove", (1,)),
        ("AddOrReplace", (1,), [6.]),
    ]
    assert mon.dropped == 2


def test_lww_equal_timestamp_does_not_modify(vst):
    from vector_store_text_b200.service import DbEmbedding, ItemsMonitor
    idx = RecordingIndex()
    mon = ItemsMonitor(idx)
    assert mon.add(DbEmbedding("a", [1.], 7))
    assert not mon.add(DbEmbedding("a", [2.], 7))     # `timestamp.0 < embedding.timestamp.0` is strict (monitor_items.rs:64)
    assert len(idx.messages) == 1


def test_engine_registry_rules(vst):
    from vector_store_text_b200.service import Engine
    made, closed = [], []

    class A:
        def __init__(self, i): self.i = i
        def close(self): closed.append(self.i)

    def factory(index_id):
        if index_id == "bad.idx":
            raise RuntimeError("unable to create")
        made.append(index_id)
        return A(index_id)

    e = Engine(factory)
    assert e.get_index_ids() == []
    assert e.add_index("ks.a") and e.get_index("ks.a").i == "ks.a"
    first = e.get_index("ks.a")
    assert not e.add_index("ks.a")                      # never replaces (engine.rs:101-104)
    assert e.get_index("ks.a") is first and made == ["ks.a"]
    assert not e.add_index("bad.idx")                   # factory error: registry unchanged (engine.rs:107-113)
    assert e.get_index_ids() == ["ks.a"]
    assert e.get_index("ks.none") is None
    e.del_index("ks.a"); e.del_index("ks.a")            # second one is a no-op
    assert e.get_index_ids() == [] and closed == ["ks.a"]


def _get(url):
    with urllib.request.urlopen(url, timeout=10) as r:
        return r.status, json.loads(r.read())


def _post(url, obj):
    req = urllib.request.Request(url, data=json.dumps(obj).encode(), headers={"Content-Type": "application/json"})
    try:
        with urllib.request.urlopen(req, timeout=10) as r:
            return r.status, r.read().decode()
    except urllib.error.HTTPError as e:
        return e.code, e.read().decode()


@pytest.mark.parametrize("native", [False, True])
def test_integration_kat_over_http(vst, oracle_lib, native):
    """tests/integration/usearch.rs:20-135 with the HTTP client of tests/integration/httpclient.rs."""
    run_integration_kat(vst, oracle_lib, native)


def run_integration_kat(vst, lib, native):
    """The KAT itself, over any backend exporting include/vsann.h (the oracle here, libvsann.so in tests/test_gpu_service.py)."""
    from vector_store_text_b200.service import DbEmbedding, Engine, HttpServer, ItemsMonitor

    def factory(index_id):
        if native:
            return vst.NativeActor(index_id, 3, metric=vst.Metric.COS, capacity_hint=100, threads=1, lib=lib)
        return vst.IndexActor(index_id, 3, metric=vst.Metric.COS, capacity_hint=100, lib=lib)

    engine = Engine(factory)
    srv = HttpServer(engine, pk_columns=lambda index_id: ["pk", "ck"])
    base = f"http://{srv.addr[0]}:{srv.addr[1]}/api/v1"
    try:
        assert _get(f"{base}/indexes") == (200, [])
        assert _post(f"{base}/indexes/vector/ann/ann", {"embedding": [1, 2, 3]})[0] == 404      # no such index yet
        with pytest.raises(urllib.error.HTTPError) as ei:
            _get(f"{base}/indexes/vector/ann/count")
        assert ei.value.code == 404

        engine.add_index("vector.ann")
        mon = ItemsMonitor(engine.get_index("vector.ann"))
        mon.feed([DbEmbedding((1, "one"), [1., 1., 1.], 10),
                  DbEmbedding((2, "two"), [2., -2., 2.], 20),
                  DbEmbedding((3, "three"), [3., 3., 3.], 30)])
        t0 = time.time()
        while _get(f"{base}/indexes/vector/ann/count")[1] != 3:       # adds are fire-and-forget: poll (usearch.rs:98-104)
            assert time.time() - t0 < 10
            time.sleep(0.005)
        assert _get(f"{base}/indexes") == (200, ["vector.ann"])

        st, body = _post(f"{base}/indexes/vector/ann/ann", {"embedding": [2.1, -2., 2.], "limit": 1})
        assert st == 200
        resp = json.loads(body)
        assert len(resp["distances"]) == 1
        assert resp["primary_keys"] == {"pk": [2], "ck": ["two"]}
        st, body = _post(f"{base}/indexes/vector/ann/ann", {"embedding": [2.1, -2., 2.]})        # limit defaults to 1
        assert st == 200 and len(json.loads(body)["distances"]) == 1
        st, body = _post(f"{base}/indexes/vector/ann/ann", {"embedding": [2.1, -2., 2.], "limit": 3})
        resp = json.loads(body)
        assert resp["primary_keys"]["pk"][0] == 2 and sorted(resp["primary_keys"]["pk"]) == [1, 2, 3]
        assert resp["distances"] == sorted(resp["distances"])

        st, body = _post(f"{base}/indexes/vector/ann/ann", {"embedding": [1., 2.], "limit": 1})  # 500 + the actor's text
        assert st == 500 and body == "index.ann request error: ann: wrong embedding dimensions: 2 != 3"
        st, body = _post(f"{base}/indexes/vector/ann/ann", {"embedding": [], "limit": 1})
        assert st == 500 and body.endswith("ann: embedding dimensions == 0")
        assert _post(f"{base}/indexes/vector/ann/ann", {"limit": 1})[0] == 422                   # body does not deserialize

        engine.del_index("vector.ann")                                                           # usearch.rs:125-134
        assert _get(f"{base}/indexes") == (200, [])
        assert _post(f"{base}/indexes/vector/ann/ann", {"embedding": [1, 2, 3]})[0] == 404
    finally:
        srv.close()
