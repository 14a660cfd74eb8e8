"""Time the gather + merge kernel on ONE GPU with every part in local memory (what is left of the exchange step without NVLink)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vector_store_text_b200 as vst
from vector_store_text_b200._ffi import merge_topk_peers_device
lib = vst.load_library(); dev = torch.device("cuda", 0)
for parts, nq, k in ((2, 10000, 10), (8, 10000, 10), (2, 4096, 100), (8, 4096, 100)):
    bufs_k = [torch.arange(nq * k, device=dev).view(nq, k) * parts + p for p in range(parts)]
    bufs_d = [torch.arange(nq * k, device=dev, dtype=torch.float32).view(nq, k) / (nq * k) + p * 1e-3 for p in range(parts)]
    kp = torch.tensor([b.data_ptr() for b in bufs_k], dtype=torch.int64, device=dev); dp = torch.tensor([b.data_ptr() for b in bufs_d], dtype=torch.int64, device=dev)
    mk = torch.empty((nq, k), dtype=torch.int64, device=dev); md = torch.empty((nq, k), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for env in ("1", "0"):
        os.environ["VSANN_MERGE_RANK"] = env
        tot = 0.0
        for step in range(25):
            torch.cuda._sleep(600000)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); merge_topk_peers_device(0, parts, nq, k, kp.data_ptr(), dp.data_ptr(), mk.data_ptr(), md.data_ptr(), stream=stream, lib=lib); e1.record(); e1.synchronize()
            if step >= 5: tot += e0.elapsed_time(e1)
        print(f"parts={parts} nq={nq} k={k} rank_merge={env}: {tot / 20 * 1e3:.1f} us", flush=True)
    # an empty kernel's worth: two events around nothing but a tiny torch op
    tot = 0.0
    for step in range(25):
        torch.cuda._sleep(600000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); mk[0, 0] = 1; e1.record(); e1.synchronize()
        if step >= 5: tot += e0.elapsed_time(e1)
    print(f"  (events around one tiny kernel: {tot / 20 * 1e3:.1f} us)", flush=True)
