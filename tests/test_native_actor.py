"""CPU tests of the C++ index actor (csrc/actor.cpp, include/vsactor.h) over the oracle backend: the message
semantics of src/index/usearch.rs:82-311 — bounded channel, 2 x T in flight, fire-and-forget adds, bimap
roll-back, replace, error strings — and the reference's own unit test (usearch.rs:323-425) through it."""
import threading
import time

import numpy as np
import pytest


def spawn(vst, oracle_lib, dims=3, metric=1, threads=2, **kw):
    return vst.NativeActor("vector.ann", dims, metric=metric, capacity_hint=1000, threads=threads, lib=oracle_lib, **kw)


def wait_count(a, n, timeout=10.0):
    """The reference's tests poll `count` until fire-and-forget adds are visible (usearch.rs:344-356)."""
    t0 = time.time()
    while time.time() - t0 < timeout:
        if a.count() == n:
            return True
        time.sleep(0.005)
    return a.count() == n


def test_library_exports_actor_abi(vst):
    import ctypes as C
    lib = C.CDLL(vst.actor_lib_path())
    for s in vst.ACTOR_SYMBOLS:
        assert hasattr(lib, s), s


@pytest.mark.parametrize("metric", [0, 1, 2])
def test_reference_unit_test_through_the_actor(vst, oracle_lib, metric):
    a = spawn(vst, oracle_lib, metric=metric)
    a.add_or_replace((1, "one"), [1, 1, 1]); a.add_or_replace((2, "two"), [2, -2, 2]); a.add_or_replace((3, "three"), [3, 3, 3])
    assert wait_count(a, 3)
    pks, d = a.ann([2.2, -2.2, 2.2], 1)
    assert pks == [(2, "two")] and len(d) == 1
    a.add_or_replace((3, "three"), [2.1, -2.1, 2.1])
    time.sleep(0.05)
    assert wait_count(a, 3)
    if metric != 0:                                   # cosine: rounding-level tie
        assert a.ann([2.2, -2.2, 2.2], 1)[0] == [(3, "three")]
    a.remove((3, "three"))
    assert wait_count(a, 2)
    assert a.ann([2.2, -2.2, 2.2], 1)[0] == [(2, "two")]
    a.remove((42, "absent"))
    assert wait_count(a, 2)
    st = a.stats()
    assert st["keys"] == 2 and st["next_key"] == 3 and st["rollbacks"] == 0    # the replace reused key 2 (usearch.rs:190-196)
    a.close()


def test_ann_error_strings(vst, oracle_lib):
    a = spawn(vst, oracle_lib)
    with pytest.raises(vst.AnnError, match="ann: embedding dimensions == 0"):
        a.ann([], 1)
    with pytest.raises(vst.AnnError, match=r"ann: wrong embedding dimensions: 2 != 3"):
        a.ann([1, 2], 1)
    with pytest.raises(ValueError):
        a.ann([1, 2, 3], 0)
    a.close()


def test_many_producers_bounded_queue_and_in_flight(vst, oracle_lib):
    """8 producer threads, 2 workers: nothing is lost, the channel never holds more than 10 messages and at most
    2 x T are in flight (usearch.rs:102,118)."""
    d, n, producers, T = 16, 4000, 8, 2
    x = np.random.default_rng(3).standard_normal((n, d)).astype(np.float32)
    a = spawn(vst, oracle_lib, dims=d, threads=T, connectivity=8, expansion_add=32)

    def produce(p):
        for i in range(p, n, producers):
            a.add_or_replace(("pk", i), x[i])

    ts = [threading.Thread(target=produce, args=(p,)) for p in range(producers)]
    [t.start() for t in ts]; [t.join() for t in ts]
    assert wait_count(a, n, 60)
    st = a.stats()
    assert st["keys"] == n and st["next_key"] == n
    assert 1 <= st["max_queued"] <= 10 and 1 <= st["max_in_flight"] <= 2 * T and st["threads"] == T
    # every vector finds itself
    hit = sum(a.ann(x[i], 1)[0] == [("pk", i)] for i in range(0, n, 97))
    assert hit >= 0.95 * len(range(0, n, 97))
    # replaces and removes keep the bimap and the index in step
    for i in range(0, 200):
        a.add_or_replace(("pk", i), x[i] + 1.0)
    for i in range(200, 300):
        a.remove(("pk", i))
    assert wait_count(a, n - 100, 60)
    assert a.stats()["keys"] == n - 100 and a.stats()["next_key"] == n
    pks, _ = a.ann(x[250], 5)
    assert ("pk", 250) not in pks
    a.close()


def test_close_drains_the_channel(vst, oracle_lib):
    a = spawn(vst, oracle_lib, dims=4)
    for i in range(50):
        a.add_or_replace(i, [i, 0, 0, 1])
    a.close()          # returns only after every queued message was processed and the index freed
    a.close()          # idempotent
