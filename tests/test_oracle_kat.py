"""CPU tests: the oracle against every numeric expectation the reference holds for this path
 and against itself (exact vs HNSW recall)."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_data, recall_at_k

V1, V2, V3 = [1, 1, 1], [2, -2, 2], [3, 3, 3]


@pytest.mark.parametrize("metric", [0, 1, 2])
def test_kat_usearch_rs_unit(oracle_lib, vst, metric):
    """src/index/usearch.rs:323-425 add_or_replace_size_ann (keys stand in for the composite PKs)."""
    ix = vst.VsIndex(3, metric=metric, capacity_hint=1000, lib=oracle_lib)
    ix.add(1, V1); ix.add(2, V2); ix.add(3, V3)
    assert ix.count() == 3
    keys, dist = ix.search([2.2, -2.2, 2.2], 1)
    assert keys.tolist() == [2] and len(dist) == 1
    ix.add(3, [2.1, -2.1, 2.1])                    # replace
    assert ix.count() == 3
    if metric != 0:                                # cosine: rounding-level tie (SURVEY.md §8c caveat)
        keys, _ = ix.search([2.2, -2.2, 2.2], 1)
        assert keys.tolist() == [3]
    assert ix.remove(3)
    assert ix.count() == 2
    keys, _ = ix.search([2.2, -2.2, 2.2], 1)
    assert keys.tolist() == [2]


@pytest.mark.parametrize("metric", [0, 1, 2])
def test_kat_integration_usearch(oracle_lib, vst, metric):
    """tests/integration/usearch.rs:74-123: query [2.1,-2.,2.] -> pk 2, exactly one distance."""
    ix = vst.VsIndex(3, metric=metric, capacity_hint=1000, lib=oracle_lib)
    for k, v in ((1, V1), (2, V2), (3, V3)):
        ix.add(k, v)
    keys, dist = ix.search([2.1, -2.0, 2.0], 1)
    assert keys.tolist() == [2] and len(dist) == 1


def test_appendix_c_distances(oracle_lib):
    """SURVEY.md Appendix C (naive sequential f32) — pins the distance definitions."""
    def d(metric, a, b):
        a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
        return oracle_lib.ovs_distance(metric, a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), 3)
    q = [2.2, -2.2, 2.2]
    assert d(1, q, V1) == pytest.approx(13.120001, rel=1e-6)
    assert d(1, q, V2) == pytest.approx(0.12000005, rel=1e-5)
    assert d(1, q, V3) == pytest.approx(28.319996, rel=1e-6)
    assert d(2, q, V1) == pytest.approx(-1.2, rel=1e-6)
    assert d(2, q, V2) == pytest.approx(-12.200001, rel=1e-6)
    assert d(0, q, V1) == pytest.approx(0.6666666, rel=1e-5)
    assert abs(d(0, q, V2)) < 3e-7
    q2 = [2.1, -2.0, 2.0]
    assert d(0, q2, V2) == pytest.approx(0.00026869774, rel=1e-3)
    assert d(1, q2, V2) == pytest.approx(0.009999981, rel=1e-4)


def test_exact_matches_numpy(oracle_lib, vst):
    x = make_data(2000, 32, 1); q = make_data(20, 32, 2)
    ix = vst.VsIndex(32, metric=1, capacity_hint=4000, lib=oracle_lib)
    ix.add_batch(np.arange(2000), x)
    keys, dist, found = ix.search(q, 10, exact=True)
    ref = ((q[:, None, :].astype(np.float64) - x[None].astype(np.float64)) ** 2).sum(-1)
    want = np.argsort(ref, axis=1, kind="stable")[:, :10]
    assert (keys == want).mean() > 0.999
    assert np.all(found == 10)
    np.testing.assert_allclose(dist, np.take_along_axis(ref, want, 1), rtol=1e-5)


@pytest.mark.parametrize("metric", [0, 1, 2])
def test_hnsw_recall_vs_exact(oracle_lib, vst, metric):
    x = make_data(5000, 32, 3, clustered=True); q = make_data(100, 32, 4, clustered=True)
    if metric == 2:   # inner product is only a proper ranking on normalised rows
        x /= np.linalg.norm(x, axis=1, keepdims=True); q /= np.linalg.norm(q, axis=1, keepdims=True)
    ix = vst.VsIndex(32, metric=metric, capacity_hint=8000, lib=oracle_lib)
    ix.add_batch(np.arange(5000), x)
    tk, _, _ = ix.search(q, 10, exact=True)
    ak, ad, _ = ix.search(q, 10, ef=64)
    assert recall_at_k(ak, tk) > 0.9
    assert np.all(np.diff(ad, axis=1) >= 0)


def test_multithreaded_build_recall(oracle_lib, vst):
    x = make_data(4000, 16, 5); q = make_data(50, 16, 6)
    ix = vst.VsIndex(16, metric=1, capacity_hint=8000, lib=oracle_lib)
    ix.set_threads(4)
    ix.add_batch(np.arange(4000), x)
    tk, _, _ = ix.search(q, 10, exact=True)
    ak, _, _ = ix.search(q, 10, ef=64)
    assert recall_at_k(ak, tk) > 0.85


def test_small_index_and_errors(oracle_lib, vst):
    ix = vst.VsIndex(4, metric=1, capacity_hint=100, lib=oracle_lib)
    keys, dist = ix.search([0, 0, 0, 0], 5)
    assert len(keys) == 0
    ix.add(7, [1, 0, 0, 0])
    keys, dist = ix.search([0, 0, 0, 0], 5)
    assert keys.tolist() == [7] and dist.tolist() == [1.0]
    assert not ix.remove(99)
    with pytest.raises(vst.VsError):
        ix.add(1, [1, 2, 3])
    with pytest.raises(vst.VsError):
        vst.VsIndex(0, lib=oracle_lib)


def test_capacity_growth(oracle_lib, vst):
    ix = vst.VsIndex(2, metric=1, capacity_hint=10, lib=oracle_lib)
    x = make_data(50, 2, 7)
    ix.add_batch(np.arange(50), x)
    assert ix.count() == 50 and ix.capacity() >= 50
