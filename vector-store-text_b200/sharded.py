"""Sharded index: one process per GPU, corpus sharded by key.

  shard(key) = key mod world_size           — a replaced key stays on its shard
  add / remove                              — routed to the owning rank, no collective
  ann                                       — every rank searches its shard with the full ef, then the
                                              exchange + G-way merge per query (K6). On GPUs the exchange
                                              is fused into the merge kernel (`PeerExchange`: result lists
                                              in symmetric memory, one stream barrier, K6 loads the peers'
                                              lists over NVLink / NVSwitch); otherwise one all_gather of
                                              (dist f32, key u64)[B x k] per rank and the dense merge
  count                                     — all_reduce(sum) of the shard counts

The collective is `torch.distributed` (NCCL over NVLink on the GPU box; gloo in the CPU tests, where
the local shard is the oracle backend and the merge below runs on host tensors — test plumbing only).
"""
import numpy as np
import torch
import torch.distributed as dist

import os

from ._ffi import VsIndex, merge_topk_device, merge_topk_flags_device, merge_topk_peers_device


class PeerExchange:
    """Fused barrier + merge over peer memory for `nq` x `k` result lists per rank.

    Every rank owns two buffers in symmetric memory and one flag word, all mapped into every peer. A buffer holds one slot of
    (keys u64 [nq, k], dists f32 [nq, k]) PER RANK. A step has the shard's lists written into slot `rank` of buffer `turn` and
    calls `merge()`: ONE launch of the merge kernel, which publishes this rank's flag = step number (release, system scope),
    waits for every peer's flag (acquire loads over NVLink, one lane per peer), and merges.
      push (default): `search()` hands the traversal kernel the address of slot `rank` in EVERY rank's buffer
        (vs_search_batch_fanout_device): a query's list goes out over NVLink as soon as the query is done, the transfer hides
        behind the rest of the traversal, and the merge reads LOCAL memory only;
      pull (`VSANN_P2P_PUSH=0`, and always for exact searches): the lists are written locally, the merge kernel loads the peers'
        slots over NVLink.
    The buffers alternate, so one flag per rank is enough: a rank that has seen a peer's flag i+1 knows that peer has finished
    merge i. `VSANN_P2P_FLAGS=0` selects torch's symmetric-memory barrier kernel followed by the merge kernel. Raises if symmetric
    memory cannot be set up (callers fall back to all_gather)."""

    def __init__(self, nq, k, device, group=None, lib=None):
        import torch.distributed._symmetric_memory as symm
        group = group or dist.group.WORLD
        self.nq, self.k, self.device, self.lib = nq, k, device, lib
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.koff = nq * k * 8
        self.slot = nq * k * 12 + (-(nq * k * 12) % 16)                      # one rank's lists: keys, then dists
        self.push = os.environ.get("VSANN_P2P_PUSH", "1") != "0" and self.world <= 9
        self.bufs = [symm.empty(self.slot * self.world, dtype=torch.uint8, device=device) for _ in range(2)]
        self.hdls = [symm.rendezvous(b, group) for b in self.bufs]
        mine = self.rank * self.slot
        # pull: part p = slot p of RANK p's buffer; push: part p = slot p of MY buffer
        self.kptr = [torch.tensor([int(p) + r * self.slot for r, p in enumerate(h.buffer_ptrs)], dtype=torch.int64, device=device) for h in self.hdls]
        self.dptr = [torch.tensor([int(p) + r * self.slot + self.koff for r, p in enumerate(h.buffer_ptrs)], dtype=torch.int64, device=device) for h in self.hdls]
        self.kptr_local = [torch.tensor([b.data_ptr() + r * self.slot for r in range(self.world)], dtype=torch.int64, device=device) for b in self.bufs]
        self.dptr_local = [torch.tensor([b.data_ptr() + r * self.slot + self.koff for r in range(self.world)], dtype=torch.int64, device=device) for b in self.bufs]
        # where MY lists go: my own buffer first (the primary output), then every peer's
        order = [self.rank] + [r for r in range(self.world) if r != self.rank]
        self.fan_keys = [[int(h.buffer_ptrs[r]) + mine for r in order] for h in self.hdls]
        self.fan_dist = [[int(h.buffer_ptrs[r]) + mine + self.koff for r in order] for h in self.hdls]
        self.flags = symm.empty(64, dtype=torch.uint8, device=device)
        self.flags.zero_()
        self.fhdl = symm.rendezvous(self.flags, group)
        self.fptr = torch.tensor([int(p) for p in self.fhdl.buffer_ptrs], dtype=torch.int64, device=device)
        self.use_flags = os.environ.get("VSANN_P2P_FLAGS", "1") != "0"
        self.epoch = 0
        self.turn = 0
        self.pushed = [False, False]
        torch.cuda.synchronize(device)
        self.fhdl.barrier(channel=0)          # every rank's flag word is zero before anybody publishes
        torch.cuda.synchronize(device)

    def describe(self):
        how = ("result lists pushed to every peer by the traversal kernel (stores over NVLink while it runs), merge from local memory"
               if self.push else "merge kernel loads the peers' lists from symmetric memory")
        if self.use_flags:
            return how + ", barrier fused into the merge kernel (release/acquire flags over NVLink): one launch after the traversal"
        return how + " (torch symmetric-memory barrier kernel per step)"

    def next_buffer(self):
        """(index, keys device pointer, dists device pointer) of this rank's slot in the buffer this step publishes into."""
        i = self.turn
        self.turn ^= 1
        self.pushed[i] = False
        base = self.bufs[i].data_ptr() + self.rank * self.slot
        return i, base, base + self.koff

    def search(self, index, nq, q_ptr, ef=0, exact=False, stream=0):
        """The shard search of a step through `index` (a VsIndex on this rank's GPU); returns the buffer index for merge()."""
        i, kp, dp = self.next_buffer()
        if self.push and not exact:
            index.search_fanout_device(nq, q_ptr, self.k, self.fan_keys[i], self.fan_dist[i], ef=ef, stream=stream)
            self.pushed[i] = True
        else:
            index.search_device(nq, q_ptr, self.k, kp, dp, ef=ef, exact=exact, stream=stream)
        return i

    def views(self, i):
        b = self.bufs[i][self.rank * self.slot:(self.rank + 1) * self.slot]
        return b[:self.koff].view(torch.int64).view(self.nq, self.k), b[self.koff:self.koff + self.nq * self.k * 4].view(torch.float32).view(self.nq, self.k)

    def merge(self, i, mk, md):
        stream = torch.cuda.current_stream().cuda_stream
        kptr, dptr = (self.kptr_local[i], self.dptr_local[i]) if self.pushed[i] else (self.kptr[i], self.dptr[i])
        if self.use_flags:
            self.epoch += 1
            merge_topk_flags_device(self.device.index, self.world, self.nq, self.k, kptr.data_ptr(), dptr.data_ptr(),
                                    self.fptr.data_ptr(), self.rank, self.epoch, mk.data_ptr(), md.data_ptr(), stream=stream, lib=self.lib)
            return mk, md
        self.hdls[i].barrier(channel=0)
        merge_topk_peers_device(self.device.index, self.world, self.nq, self.k, kptr.data_ptr(), dptr.data_ptr(),
                                mk.data_ptr(), md.data_ptr(), stream=stream, lib=self.lib)
        return mk, md


def shard_of(keys, world):
    return np.asarray(keys, dtype=np.uint64) % np.uint64(world)


def merge_topk_host(keys, dists, k):
    """Reference merge on host tensors: keys/dists [G, nq, k] ascending per part -> [nq, k] by (dist, key).
    Used by the gloo tests and as the checker of the CUDA merge kernel."""
    g, nq, _ = dists.shape
    d = dists.permute(1, 0, 2).reshape(nq, -1).numpy()
    kk = keys.permute(1, 0, 2).reshape(nq, -1).numpy().view(np.uint64)
    order = np.lexsort((kk, d), axis=1)[:, :k]
    return torch.from_numpy(np.take_along_axis(kk, order, 1).view(np.int64).copy()), torch.from_numpy(np.take_along_axis(d, order, 1).copy())


class ShardedIndex:
    def __init__(self, dims, rank=None, world=None, device=None, lib=None, **opts):
        self.rank = dist.get_rank() if rank is None else rank
        self.world = dist.get_world_size() if world is None else world
        self.cuda = device is not None
        self.device = device
        self.local = VsIndex(dims, device=(device.index if self.cuda else 0), lib=lib, **opts)
        self.dims = dims
        self._lib = lib
        self._px = {}

    # -- mutation: each rank is handed the full batch and keeps its own keys -----------------------
    def add_batch(self, keys, vecs):
        keys = np.asarray(keys, dtype=np.uint64)
        mine = shard_of(keys, self.world) == np.uint64(self.rank)
        if mine.any():
            self.local.add_batch(keys[mine], np.asarray(vecs, dtype=np.float32)[mine])

    def remove(self, key):
        if int(key) % self.world == self.rank:
            return self.local.remove(key)
        return False

    def count(self):
        t = torch.tensor([self.local.count()], dtype=torch.int64, device=self.device if self.cuda else "cpu")
        dist.all_reduce(t)
        return int(t.item())

    # -- search ---------------------------------------------------------------------------------------
    def ann(self, queries, k, ef=0, exact=False):
        """queries: [nq, dims] float32 (identical on every rank). Returns (keys u64 [nq,k], dists [nq,k])."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        nq = q.shape[0]
        if self.cuda:
            qd = torch.from_numpy(q).to(self.device)
            ok = torch.empty((nq, k), dtype=torch.int64, device=self.device)
            od = torch.empty((nq, k), dtype=torch.float32, device=self.device)
            mk, md = self.ann_device(qd, ok, od, k, ef, exact)
            return mk.cpu().numpy().view(np.uint64), md.cpu().numpy()
        lk, ld, _ = self.local.search(q, k, ef=ef, exact=exact)
        ok = torch.from_numpy(lk.view(np.int64).copy()); od = torch.from_numpy(ld.copy())
        gk = torch.empty((self.world, nq, k), dtype=torch.int64); gd = torch.empty((self.world, nq, k), dtype=torch.float32)
        dist.all_gather_into_tensor(gk.view(-1), ok.view(-1))
        dist.all_gather_into_tensor(gd.view(-1), od.view(-1))
        mk, md = merge_topk_host(gk, gd, k)
        return mk.numpy().view(np.uint64), md.numpy()

    def peer_exchange(self, nq, k):
        """The PeerExchange for this (nq, k) shape, or None when symmetric memory is unavailable (collective: every rank calls)."""
        key = (nq, k)
        if key not in self._px:
            try:
                self._px[key] = PeerExchange(nq, k, self.device, lib=self._lib)
            except Exception:   # noqa: BLE001
                self._px[key] = None
        return self._px[key]

    def ann_device(self, qd, ok, od, k, ef=0, exact=False, gk=None, gd=None, mk=None, md=None, exchange="auto"):
        """Device-resident path on the current torch stream: shard search -> exchange -> merge kernel.
        exchange: "auto" (peer memory when it comes up, else all_gather), "p2p", "all_gather"."""
        nq = qd.shape[0]
        stream = torch.cuda.current_stream().cuda_stream
        if self.world > 1 and exchange != "all_gather":
            px = self.peer_exchange(nq, k)
            if px is None and exchange == "p2p":
                raise RuntimeError("symmetric memory is not available")
            if px is not None:
                i = px.search(self.local, nq, qd.data_ptr(), ef=ef, exact=exact, stream=stream)
                mk = torch.empty_like(ok) if mk is None else mk
                md = torch.empty_like(od) if md is None else md
                return px.merge(i, mk, md)
        self.local.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), ef=ef, exact=exact, stream=stream)
        if self.world == 1:
            return ok, od
        gk = torch.empty((self.world, nq, k), dtype=torch.int64, device=self.device) if gk is None else gk
        gd = torch.empty((self.world, nq, k), dtype=torch.float32, device=self.device) if gd is None else gd
        mk = torch.empty_like(ok) if mk is None else mk
        md = torch.empty_like(od) if md is None else md
        dist.all_gather_into_tensor(gk.view(-1), ok.view(-1))
        dist.all_gather_into_tensor(gd.view(-1), od.view(-1))
        merge_topk_device(self.device.index, self.world, nq, k, gk.data_ptr(), gd.data_ptr(), mk.data_ptr(), md.data_ptr(), stream=stream, lib=self._lib)
        return mk, md

    def close(self):
        self.local.close()
