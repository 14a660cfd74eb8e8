"""Index persistence (include/vsann.h vs_index_save / vs_index_load). The reference has none
(it rebuilds from a table scan, src/db_index.rs:108-117,217-237), so the bar is self-consistency: a restored index
answers exactly like the saved one and keeps accepting add / replace / remove."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import make_data


def test_load_rejects_bad_files_without_a_gpu(vst, tmp_path):
    """Header validation happens before any CUDA call, so these run on the CPU box too."""
    lib = vst.load_library()
    h = C.c_void_p()
    missing = os.fsencode(str(tmp_path / "nope.vsann"))
    assert lib.vs_index_load(missing, None, C.byref(h)) == vst.VS_ERR_INVALID
    assert b"cannot open" in lib.vs_last_error()
    junk = tmp_path / "junk.vsann"
    junk.write_bytes(b"not an index" * 50)
    assert lib.vs_index_load(os.fsencode(str(junk)), None, C.byref(h)) == vst.VS_ERR_INVALID
    assert b"not a vsann index file" in lib.vs_last_error()
    newer = tmp_path / "newer.vsann"
    newer.write_bytes(b"VSANNIDX" + (99).to_bytes(4, "little") + bytes(116))
    assert lib.vs_index_load(os.fsencode(str(newer)), None, C.byref(h)) == vst.VS_ERR_UNSUPPORTED
    corrupt = tmp_path / "corrupt.vsann"
    corrupt.write_bytes(b"VSANNIDX" + (1).to_bytes(4, "little") + bytes(116))     # dims == 0
    assert lib.vs_index_load(os.fsencode(str(corrupt)), None, C.byref(h)) == vst.VS_ERR_INVALID
    assert b"corrupt header" in lib.vs_last_error()
    assert lib.vs_index_save(None, missing) == vst.VS_ERR_INVALID


@pytest.mark.gpu
@pytest.mark.parametrize("metric,d", [(1, 128), (0, 100), (2, 768)])
def test_round_trip_answers_identically(vst, tmp_path, metric, d):
    n = 20000 if d <= 128 else 6000
    x = make_data(n, d, 5, clustered=True); q = make_data(200, d, 6, clustered=True)
    keys = (np.arange(n, dtype=np.uint64) * 7 + 3)
    g = vst.VsIndex(d, metric=vst.Metric(metric), capacity_hint=n)
    g.add_batch(keys, x)
    g.add_batch(keys[:500], x[:500] * 1.5 + 0.25)          # replaces: tombstones + new slots
    for k in keys[1000:1100]:
        assert g.remove(int(k))
    g.add(10**12, x[0])                                     # staged, not yet landed: save flushes it
    path = str(tmp_path / "ix.vsann")
    g.save(path)
    a_k, a_d, a_f = g.search(q, 10, ef=64)
    e_k, e_d, e_f = g.search(q, 10, exact=True)
    st = g.stats()

    r = vst.VsIndex.load(path, capacity_hint=n)
    assert r.dims == d and r.count() == g.count() == n - 100 + 1
    st2 = r.stats()
    assert (st2["slots"], st2["top_level"]) == (st["slots"], st["top_level"])
    b_k, b_d, b_f = r.search(q, 10, ef=64)
    assert np.array_equal(a_k, b_k) and np.array_equal(a_d, b_d) and np.array_equal(a_f, b_f)
    x_k, x_d, x_f = r.search(q, 10, exact=True)
    assert np.array_equal(e_k, x_k) and np.array_equal(e_d, x_d)
    for lvl in range(st["top_level"] + 1):                  # every link table, bit for bit
        la, lb = g.export_level(lvl), r.export_level(lvl)
        assert np.array_equal(la[0], lb[0]) and np.array_equal(la[1], lb[1])

    # the restored index keeps working: same mutations on both give the same answers
    extra = make_data(3000, d, 7, clustered=True)
    ek = np.arange(3000, dtype=np.uint64) + 10**9
    for ix in (g, r):
        ix.add_batch(ek, extra)
        ix.add_batch(keys[2000:2100], extra[:100])          # replace keys that were live in the file
        assert ix.remove(int(keys[5])) and not ix.remove(int(keys[1000]))   # removed before the save: still unknown
        ix.flush()
    assert r.count() == g.count()
    c_k, c_d, _ = g.search(q, 10, exact=True)
    d_k, d_d, _ = r.search(q, 10, exact=True)
    assert np.array_equal(c_k, d_k) and np.array_equal(c_d, d_d)
    f_k, _, _ = g.search(q, 10, ef=64)
    h_k, _, _ = r.search(q, 10, ef=64)
    assert np.array_equal(f_k, h_k)
    g.close(); r.close()


@pytest.mark.gpu
def test_empty_index_and_damaged_file(vst, tmp_path):
    g = vst.VsIndex(16, metric=vst.Metric.L2SQ, capacity_hint=1000)
    p0 = str(tmp_path / "empty.vsann")
    g.save(p0)
    r = vst.VsIndex.load(p0, capacity_hint=1000)
    assert r.count() == 0 and r.dims == 16
    r.add_batch(np.arange(50, dtype=np.uint64), make_data(50, 16, 1))
    assert r.count() == 50
    x = make_data(5000, 16, 2)
    g.add_batch(np.arange(5000, dtype=np.uint64), x)
    p1 = str(tmp_path / "full.vsann")
    g.save(p1)
    raw = bytearray(open(p1, "rb").read())
    raw[len(raw) // 2] ^= 0x40                               # one flipped bit in the payload
    bad = str(tmp_path / "flipped.vsann"); open(bad, "wb").write(raw)
    with pytest.raises(vst.VsError, match="checksum mismatch"):
        vst.VsIndex.load(bad)
    cut = str(tmp_path / "cut.vsann"); open(cut, "wb").write(raw[: len(raw) // 3])
    with pytest.raises(vst.VsError, match="truncated"):
        vst.VsIndex.load(cut)
    g.close(); r.close()
