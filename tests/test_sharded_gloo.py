"""Multi-shard host logic on CPU: world_size-2 gloo processes, each holding the shard
`key % world` behind the oracle backend (test plumbing: the GPU library has no CPU path). Checks that
routing, the all_gather + top-k merge and the all_reduce count give exactly what one unsharded exact
index gives."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    import vector_store_text_b200 as vst
    from vector_store_text_b200.sharded import ShardedIndex
    from conftest import make_data
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        olib = vst.load_library(os.path.join(ROOT, "oracle", "liboracle_vsann.so"), prefix="ovs_")
        n, d, nq, k = 3000, 24, 64, 10
        x = make_data(n, d, 5, clustered=True)
        q = make_data(nq, d, 6, clustered=True)
        keys = np.arange(n, dtype=np.uint64) * 7 + 3
        sh = ShardedIndex(d, metric=1, capacity_hint=n, lib=olib)
        sh.add_batch(keys, x)
        assert sh.count() == n
        sk, sd = sh.ann(q, k, exact=True)
        # single unsharded oracle index = ground truth of the merge
        full = vst.VsIndex(d, metric=1, capacity_hint=n, lib=olib)
        full.add_batch(keys, x)
        fk, fd, _ = full.search(q, k, exact=True)
        assert np.array_equal(sk, fk), "sharded exact top-k must equal the unsharded one"
        assert np.array_equal(sd.view(np.uint32), fd.view(np.uint32))
        # ANN: every shard is searched at full width => recall of the merge >= 0.9 here
        ak, ad = sh.ann(q, k, ef=64)
        rec = sum(len(set(a.tolist()) & set(t.tolist())) for a, t in zip(ak, fk)) / fk.size
        assert rec > 0.9, rec
        assert np.all(np.diff(ad, axis=1) >= 0)
        # removes are routed to the owning shard only; a replaced key stays on its shard
        for kk in keys[:50]:
            sh.remove(int(kk))
        assert sh.count() == n - 50
        sh.add_batch(keys[100:110], x[200:210])
        assert sh.count() == n - 50
        rk, _ = sh.ann(q, k, exact=True)
        assert not np.isin(rk, keys[:50]).any()
        # fewer live vectors than k on some shard: sentinel padding must lose the merge
        tiny = ShardedIndex(d, metric=1, capacity_hint=16, lib=olib)
        tiny.add_batch(keys[:3], x[:3])
        tk, td = tiny.ann(q[:4], 5, exact=True)
        assert np.all(tk[:, :3] != np.uint64(0xFFFFFFFFFFFFFFFF)) and np.all(tk[:, 3:] == np.uint64(0xFFFFFFFFFFFFFFFF))
        assert np.all(np.isinf(td[:, 3:]))
        ret[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        import traceback
        ret[rank] = "".join(traceback.format_exception(e))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_index_world2_gloo(oracle_lib):
    world = 2
    ctx = mp.get_context("spawn")
    with ctx.Manager() as man:
        ret = man.dict()
        port = _free_port()
        ps = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
        for p in ps:
            p.start()
        for p in ps:
            p.join(280)
        for p in ps:
            if p.is_alive():
                p.kill()
        out = dict(ret)
    assert out == {0: "ok", 1: "ok"}, out


def test_merge_topk_host_matches_bruteforce():
    import torch
    sys.path.insert(0, ROOT)
    import vector_store_text_b200  # noqa: F401
    from vector_store_text_b200.sharded import merge_topk_host, shard_of
    rng = np.random.default_rng(3)
    g, nq, k = 4, 33, 7
    d = np.sort(rng.integers(0, 20, (g, nq, k)).astype(np.float32), axis=2)     # many ties
    keys = rng.permutation(g * nq * k).astype(np.uint64).reshape(g, nq, k)
    mk, md = merge_topk_host(torch.from_numpy(keys.view(np.int64)), torch.from_numpy(d), k)
    for i in range(nq):
        want = sorted(zip(d[:, i].ravel().tolist(), keys[:, i].ravel().tolist()))[:k]
        assert md[i].tolist() == [w[0] for w in want]
        assert mk[i].numpy().view(np.uint64).tolist() == [w[1] for w in want]
    assert shard_of([0, 1, 2, 3, 10], 4).tolist() == [0, 1, 2, 3, 2]
