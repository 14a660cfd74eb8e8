"""PeerExchange on REAL peers (run under torchrun, one rank per GPU): every step each rank publishes random ascending (dist, key) lists
into its symmetric buffer, the fused barrier + gather + merge kernel (flags over NVLink) merges them, and the result is compared with
the host reference merge of the all-gathered lists. Ranks are skewed on purpose (random sleeps) so that a missing acquire / release
or a buffer reused too early shows up as a mismatch.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/peer_check.py"""
import os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vector_store_text_b200 as vst
from vector_store_text_b200.sharded import PeerExchange, merge_topk_host

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = vst.load_library()
bad = 0
for flags in ("1", "0"):
    os.environ["VSANN_P2P_FLAGS"] = flags
    for nq, k in ((10000, 10), (513, 100), (1, 10)):
        px = PeerExchange(nq, k, dev, lib=lib)
        mk = torch.empty((nq, k), dtype=torch.int64, device=dev); md = torch.empty((nq, k), dtype=torch.float32, device=dev)
        rng = np.random.default_rng(1000 * rank + nq)
        for step in range(25):
            d = np.sort(rng.random((nq, k), dtype=np.float32), axis=1)
            d[:, -1][rng.random(nq) < 0.1] = np.inf                      # short lists: padded tail
            keys = (rng.integers(0, 1 << 40, (nq, k)).astype(np.int64) * world + rank)
            keys[np.isinf(d)] = -1                                       # UINT64_MAX
            o = np.lexsort((keys.view(np.uint64), d), axis=1)            # lists are ascending by (distance, key), as the shard searches emit them
            d = np.take_along_axis(d, o, 1); keys = np.take_along_axis(keys, o, 1)
            i, kp, dp = px.next_buffer()
            vk, vd = px.views(i)
            if rng.random() < 0.5:
                time.sleep(float(rng.random()) * 0.003)                 # skew the ranks
            vk.copy_(torch.from_numpy(keys)); vd.copy_(torch.from_numpy(d))
            px.merge(i, mk, md)
            gk = torch.empty((world, nq, k), dtype=torch.int64, device=dev); gd = torch.empty((world, nq, k), dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(gk.view(-1), torch.from_numpy(keys).to(dev).view(-1)); dist.all_gather_into_tensor(gd.view(-1), torch.from_numpy(d).to(dev).view(-1))
            torch.cuda.synchronize()
            wk, wd = merge_topk_host(gk.cpu(), gd.cpu(), k)
            ok = np.array_equal(mk.cpu().numpy().view(np.uint64), wk.numpy().view(np.uint64)) and np.array_equal(md.cpu().numpy(), wd.numpy())
            bad += not ok
        if rank == 0:
            print(f"flags={flags} nq={nq} k={k}: 25 steps, mismatches so far {bad} ({px.describe()[:60]}...)", flush=True)
# the whole sharded step on real peers: every rank searches its own small index, the traversal kernel PUSHES its lists into every
# peer's buffer (vs_search_batch_fanout_device), the merge reads local memory - against all_gather + host merge of the same searches
n_loc, d_, nq_, k_ = 20000, 32, 2000, 10
rng = np.random.default_rng(77 + rank)
xs = rng.standard_normal((n_loc, d_), dtype=np.float32)
g = vst.VsIndex(d_, metric=1, capacity_hint=n_loc, device=local)
g.add_batch((np.arange(n_loc, dtype=np.uint64) * world + rank), xs)
g.flush()
for push in ("1", "0"):
    os.environ["VSANN_P2P_PUSH"] = push; os.environ["VSANN_P2P_FLAGS"] = "1"
    px = PeerExchange(nq_, k_, dev, lib=lib)
    mk = torch.empty((nq_, k_), dtype=torch.int64, device=dev); md = torch.empty((nq_, k_), dtype=torch.float32, device=dev)
    ok_ = torch.empty((nq_, k_), dtype=torch.int64, device=dev); od_ = torch.empty((nq_, k_), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for step in range(12):
        qs = torch.from_numpy(np.random.default_rng(500 + step).standard_normal((nq_, d_), dtype=np.float32)).to(dev)   # the same queries on every rank
        if (step + rank) % 3 == 0:
            time.sleep(0.002)
        i = px.search(g, nq_, qs.data_ptr(), ef=24, stream=stream)
        px.merge(i, mk, md)
        g.search_device(nq_, qs.data_ptr(), k_, ok_.data_ptr(), od_.data_ptr(), ef=24, stream=stream)
        gk = torch.empty((world, nq_, k_), dtype=torch.int64, device=dev); gd = torch.empty((world, nq_, k_), dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        dist.all_gather_into_tensor(gk.view(-1), ok_.view(-1)); dist.all_gather_into_tensor(gd.view(-1), od_.view(-1))
        torch.cuda.synchronize()
        wk, wd = merge_topk_host(gk.cpu(), gd.cpu(), k_)
        okk = np.array_equal(mk.cpu().numpy().view(np.uint64), wk.numpy().view(np.uint64)) and np.array_equal(md.cpu().numpy(), wd.numpy())
        bad += not okk
    if rank == 0:
        print(f"sharded step push={push}: 12 steps, mismatches so far {bad} ({px.describe()[:70]}...)", flush=True)
os.environ.pop("VSANN_P2P_PUSH", None)
g.close()

# what the exchange step costs when every peer's lists are ready: starts aligned by a device barrier, events around the one merge launch
os.environ["VSANN_P2P_FLAGS"] = "1"
for nq, k in ((10000, 10), (4096, 100)):
    px = PeerExchange(nq, k, dev, lib=lib)
    mk = torch.empty((nq, k), dtype=torch.int64, device=dev); md = torch.empty((nq, k), dtype=torch.float32, device=dev)
    for i in range(2):
        vk, vd = px.views(i)
        vk.copy_(torch.arange(nq * k, device=dev).view(nq, k) * world + rank); vd.copy_(torch.arange(nq * k, device=dev, dtype=torch.float32).view(nq, k) / (nq * k) + rank * 1e-3)
    from vector_store_text_b200._ffi import merge_topk_peers_device
    stream = torch.cuda.current_stream().cuda_stream
    def variant(name, i):
        if name == "fused flags": px.merge(i, mk, md)
        elif name == "gather + merge only": merge_topk_peers_device(dev.index, world, nq, k, px.kptr[i].data_ptr(), px.dptr[i].data_ptr(), mk.data_ptr(), md.data_ptr(), stream=stream, lib=lib)
        elif name == "torch barrier only": px.hdls[i].barrier(channel=0)
    for name in ("fused flags", "gather + merge only", "torch barrier only"):
        tot, n_it = 0.0, 40
        for step in range(n_it + 5):
            i, kp, dp = px.next_buffer()
            dist.barrier(); px.fhdl.barrier(channel=1)
            torch.cuda._sleep(600000)             # ~0.3 ms of GPU time: the launches below are queued behind it, no host latency between the events
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); variant(name, i); e1.record(); e1.synchronize()
            if step >= 5: tot += e0.elapsed_time(e1)
        t = torch.tensor([tot / n_it], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"exchange step alone (peers ready, launches queued) nq={nq} k={k} [{name}]: {t.item() * 1e3:.1f} us, max over {world} ranks", flush=True)
t = torch.tensor([bad], device=dev); dist.all_reduce(t)
if rank == 0:
    print("PEER CHECK", "PASS" if int(t.item()) == 0 else f"FAIL ({int(t.item())} mismatching steps)", f"on {world} ranks", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 0 else 1)
