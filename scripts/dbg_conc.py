"""debug: which concurrent actor breaks the graph in test_concurrent_add_search_remove"""
import sys, os, threading, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import vector_store_text_b200 as vst
from conftest import make_data, recall_at_k
n0, n1, d = 50000, 400000, 64
x = make_data(n0 + n1 + 5000, d, 81, clustered=True); q = make_data(64, d, 82, clustered=True)

def run(who):
    g = vst.VsIndex(d, metric=1, capacity_hint=100000)
    g.add_batch(np.arange(n0, dtype=np.uint64), x[:n0])
    stop = threading.Event()
    def bulk():
        g.add_batch(np.arange(n0, n0 + n1, dtype=np.uint64), x[n0:n0 + n1]); stop.set()
    def trickle():
        i = n0 + n1
        while not stop.is_set() and i < n0 + n1 + 5000:
            g.add(i, x[i]); i += 1
    def remover():
        kk = 0
        while not stop.is_set() and kk < n0:
            g.remove(kk); kk += 7
    def searcher(tid):
        while not stop.is_set():
            if tid % 2: g.search_one(q[tid % 64], 10)
            else: g.search(q[:8], 10, ef=32)
    ths = [threading.Thread(target=bulk)]
    if "t" in who: ths.append(threading.Thread(target=trickle))
    if "r" in who: ths.append(threading.Thread(target=remover))
    if "s" in who: ths += [threading.Thread(target=searcher, args=(i,)) for i in range(13)]
    if "1" in who: ths += [threading.Thread(target=searcher, args=(1,))]
    if "2" in who: ths += [threading.Thread(target=searcher, args=(2,))]
    for t in ths: t.start()
    for t in ths: t.join()
    g.flush()
    truth, _, _ = g.search(q, 10, exact=True)
    r = recall_at_k(g.search(q, 10, ef=64)[0], truth)
    st = g.stats()
    print(f"who={who!r:8} recall {r:.4f} count {st['count']} slots {st['slots']}", flush=True)
    g.close()

for who in sys.argv[1:] or ["", "t", "r", "s", "1", "2", "trs"]:
    run(who)
