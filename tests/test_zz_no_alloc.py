"""vs_exact_search_noalloc_device (include/vsann.h; no counterpart in the reference):
the allocation-free brute force, through the C ABI, against the existing exact entry point vs_search_batch(exact = 1) on the same
rows - ids and distance bits identical (the tolerance of every exact-mode test here), on two calls in a row that share the static
scratch, plus the size limits. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_data

pytestmark = pytest.mark.gpu


def noalloc(lib, torch, metric, xd, qd, k, keys_d=None, stream=None):
    n, d = xd.shape
    nq = qd.shape[0]
    ko = torch.full((nq, k), -7, dtype=torch.int64, device=xd.device)       # poisoned outputs: every entry must be written
    do = torch.full((nq, k), -7.0, dtype=torch.float32, device=xd.device)
    st = lib.vs_exact_search_noalloc_device(xd.device.index, int(metric), d, n, C.c_void_p(xd.data_ptr()),
                                            C.c_void_p(keys_d.data_ptr()) if keys_d is not None else None, nq, C.c_void_p(qd.data_ptr()), k,
                                            C.c_void_p(ko.data_ptr()), C.c_void_p(do.data_ptr()), C.c_void_p(stream) if stream else None)
    return st, ko, do


def reference(vst, metric, x, keys, q, k):
    """The existing entry point: an index over the same rows, exact mode."""
    g = vst.VsIndex(x.shape[1], metric=metric, capacity_hint=len(x), connectivity=2, expansion_add=8)   # exact search ignores the graph
    g.add_batch(keys, x)
    kk, dd, found = g.search(q, k, exact=True)
    g.close()
    return kk, dd, found


def test_noalloc_exact_matches_the_existing_entry_point_twice_in_a_row(vst):
    import torch
    lib = vst.load_library()
    dev = torch.device("cuda", 0)
    # call 1 fills most of the scratch; call 2 is smaller with another metric, k and dims: stale lists of call 1 lie right behind its own
    cases = [(1, 30000, 96, 250, 40), (0, 7001, 33, 37, 5)]       # 250 x roundup32(40) = 16000 of the 16384 the scratch holds
    outs = []
    for metric, n, d, nq, k in cases:
        x = make_data(n, d, 900 + n); q = make_data(nq, d, 901 + n)
        x[n // 2] = x[n // 3]                                            # an exact tie: the key order must decide
        q[0] = x[n // 3]
        xd, qd = torch.from_numpy(x).to(dev), torch.from_numpy(q).to(dev)
        st, ko, do = noalloc(lib, torch, metric, xd, qd, k)
        assert st == 0, lib.vs_last_error().decode()
        torch.cuda.synchronize()
        outs.append((metric, x, q, k, ko.cpu().numpy().view(np.uint64), do.cpu().numpy()))
    for metric, x, q, k, gk, gd in outs:                                # compare after BOTH calls ran back to back
        wk, wd, found = reference(vst, metric, x, np.arange(len(x), dtype=np.uint64), q, k)
        assert np.all(found == k)
        assert np.array_equal(gk, wk), "ids must equal the existing exact entry point's"
        assert np.array_equal(gd.view(np.uint32), wd.view(np.uint32)), "distances must carry the same bits"


def test_noalloc_keys_stream_short_corpus_and_no_allocation(vst):
    import torch
    lib = vst.load_library()
    dev = torch.device("cuda", 0)
    n, d, nq, k = 5000, 64, 64, 16
    x = make_data(n, d, 77); q = make_data(nq, d, 78)
    keys = (np.random.default_rng(5).permutation(n).astype(np.uint64) * 3 + 11)   # key order != row order
    x[10] = x[4000]; q[3] = x[10]
    xd, qd, kd = torch.from_numpy(x).to(dev), torch.from_numpy(q).to(dev), torch.from_numpy(keys.view(np.int64)).to(dev)
    s = torch.cuda.Stream(device=dev)
    s.wait_stream(torch.cuda.current_stream())
    st, ko, do = noalloc(lib, torch, 2, xd, qd, k, keys_d=kd, stream=s.cuda_stream)
    assert st == 0, lib.vs_last_error().decode()
    s.synchronize()
    wk, wd, _ = reference(vst, 2, x, keys, q, k)
    assert np.array_equal(ko.cpu().numpy().view(np.uint64), wk) and np.array_equal(do.cpu().numpy().view(np.uint32), wd.view(np.uint32))
    # fewer rows than k: the tail is padded
    st, ko, do = noalloc(lib, torch, 1, xd[:5].contiguous(), qd[:4].contiguous(), 8)
    assert st == 0
    ko, do = ko.cpu().numpy().view(np.uint64), do.cpu().numpy()
    assert np.all(ko[:, 5:] == np.iinfo(np.uint64).max) and np.all(np.isinf(do[:, 5:])) and np.all(np.sort(ko[:, :5], axis=1) == np.arange(5))
    # a repeated call of a shape already seen (kernels loaded, address known) leaves the device's free memory untouched
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    st, ko2, do2 = None, None, None
    ko2 = torch.empty((nq, k), dtype=torch.int64, device=dev); do2 = torch.empty((nq, k), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)                              # after the CALLER's own output buffers
    st = lib.vs_exact_search_noalloc_device(0, 2, d, n, C.c_void_p(xd.data_ptr()), C.c_void_p(kd.data_ptr()), nq, C.c_void_p(qd.data_ptr()), k,
                                            C.c_void_p(ko2.data_ptr()), C.c_void_p(do2.data_ptr()), None)
    assert st == 0
    free2, _ = torch.cuda.mem_get_info(dev)
    assert free2 == free1, (free0, free1, free2)
    assert np.array_equal(ko2.cpu().numpy().view(np.uint64), wk)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(vst):
    import torch
    lib = vst.load_library()
    dev = torch.device("cuda", 0)
    xd = torch.zeros((64, 8), device=dev); qd = torch.zeros((1024, 8), device=dev)
    st, _, _ = noalloc(lib, torch, 1, xd, qd, 32)                        # 1024 x 32 = 32768 > 16384
    assert st == vst.VS_ERR_UNSUPPORTED and "static scratch limit of 16384" in lib.vs_last_error().decode()
    st, _, _ = noalloc(lib, torch, 1, xd, qd[:4].contiguous(), 300)      # k > 256
    assert st == vst.VS_ERR_UNSUPPORTED and "k must be in [1, 256]" in lib.vs_last_error().decode()
    st, _, _ = noalloc(lib, torch, 1, torch.zeros((4, 5000), device=dev), torch.zeros((2, 5000), device=dev), 2)
    assert st == vst.VS_ERR_DIMS
    st, _, _ = noalloc(lib, torch, 1, xd, qd[:512].contiguous(), 32)     # exactly at the limit
    assert st == 0
