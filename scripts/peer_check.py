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
t = torch.tensor([bad], device=dev); dist.all_reduce(t)
if rank == 0:
    print("PEER CHECK", "PASS" if int(t.item()) == 0 else f"FAIL ({int(t.item())} mismatching steps)", f"on {world} ranks", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 0 else 1)
