"""Where the end-to-end step (vs_search_batch with pinned host buffers) spends its time beyond the traversal kernel."""
import ctypes as C, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vector_store_text_b200 as vst
n, nq, k, ef = int(os.environ.get("N", 1000000)), 10000, 10, int(os.environ.get("EF", 20))
x = bench.gen(n, 128, 42, "clustered"); q = bench.gen(nq, 128, 43, "clustered")
xd = torch.from_numpy(x).cuda()
lib = vst.load_library()
g = vst.VsIndex(128, metric=vst.Metric.L2SQ, connectivity=16, expansion_add=128, expansion_search=64, capacity_hint=n, lib=lib)
g.add_batch_device(np.arange(n, dtype=np.uint64), xd.data_ptr(), n)
qd = torch.from_numpy(q).cuda(); ok = torch.empty((nq, k), dtype=torch.int64, device="cuda"); od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
def dev():
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.search_device(nq, qd.data_ptr(), k, ok.data_ptr(), od.data_ptr(), ef=ef, stream=s); e1.record(); e1.synchronize()
    return e0.elapsed_time(e1)
for _ in range(3): dev()
print(f"device-resident kernel: {np.mean([dev() for _ in range(10)]):.3f} ms")
qh = torch.from_numpy(q).pin_memory(); kh = torch.empty((nq, k), dtype=torch.int64).pin_memory(); dh = torch.empty((nq, k), dtype=torch.float32).pin_memory()
fh = np.zeros(nq, dtype=np.uintp)
def host(found=True, qbuf=qh):
    t0 = time.perf_counter()
    st = lib.vs_search_batch(g._h, nq, C.c_void_p(qbuf.data_ptr()), k, ef, 0, C.c_void_p(kh.data_ptr()), C.c_void_p(dh.data_ptr()), fh.ctypes.data_as(C.c_void_p) if found else None)
    assert st == 0
    return (time.perf_counter() - t0) * 1e3
for name, kw in (("pinned, found[]", {}), ("pinned, found=NULL", {"found": False})):
    for _ in range(3): host(**kw)
    print(f"vs_search_batch {name}: {np.mean([host(**kw) for _ in range(20)]):.3f} ms")
qp = torch.from_numpy(q.copy())
for _ in range(3): host(qbuf=qp)
print(f"vs_search_batch pageable queries: {np.mean([host(qbuf=qp) for _ in range(20)]):.3f} ms")
