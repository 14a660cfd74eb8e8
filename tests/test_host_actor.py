"""CPU tests of the host-side mirror of the index actor (bimap, key allocation, error strings),
run over the oracle backend so no GPU is needed. Mirrors src/index/usearch.rs:313-426."""
import pytest


def make_actor(vst, oracle_lib, metric=1):
    return vst.IndexActor("vector.ann", 3, metric=metric, capacity_hint=1000, lib=oracle_lib)


def test_add_or_replace_size_ann(vst, oracle_lib):
    a = make_actor(vst, oracle_lib)
    a.add_or_replace((1, "one"), [1, 1, 1])
    a.add_or_replace((2, "two"), [2, -2, 2])
    a.add_or_replace((3, "three"), [3, 3, 3])
    assert a.count() == 3
    pks, d = a.ann([2.2, -2.2, 2.2], 1)
    assert pks == [(2, "two")] and len(d) == 1
    a.add_or_replace((3, "three"), [2.1, -2.1, 2.1])
    assert a.count() == 3
    assert a.ann([2.2, -2.2, 2.2], 1)[0] == [(3, "three")]
    a.remove((3, "three"))
    assert a.count() == 2
    assert a.ann([2.2, -2.2, 2.2], 1)[0] == [(2, "two")]
    a.remove((42, "absent"))          # unknown PK: silently ignored (usearch.rs:240-242)
    assert a.count() == 2


def test_ann_error_strings(vst, oracle_lib):
    a = make_actor(vst, oracle_lib)
    with pytest.raises(vst.AnnError, match="ann: embedding dimensions == 0"):
        a.ann([], 1)
    with pytest.raises(vst.AnnError, match=r"ann: wrong embedding dimensions: 2 != 3"):
        a.ann([1, 2], 1)


def test_failed_add_rolls_bimap_back(vst, oracle_lib):
    a = make_actor(vst, oracle_lib)
    a.add_or_replace("bad", [1, 2])           # wrong dims: add fails, bimap entry removed (usearch.rs:230-232)
    assert a.count() == 0
    assert "bad" not in a._pk_to_key
    a.add_or_replace("ok", [1, 2, 3])
    assert a.count() == 1


def test_factory(vst, oracle_lib):
    f = vst.IndexFactory(metric=vst.Metric.L2SQ, capacity_hint=100, lib=oracle_lib)
    a = f.create_index("ks.idx", 3)
    a.add_or_replace("p", [0, 0, 1])
    assert a.ann([0, 0, 1], 5)[0] == ["p"]
