#!/bin/bash
# GPU call r02k2: the wide-beam changes with the size rule for the tag hash, in-tree library: tests, then the two sweeps again.
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r02k2_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02k2_pytest.log
VSANN_BENCH_SWEEP=1 timeout 600 python bench.py --config c4 --n 12500000 --no-cpu --no-extra --steps 6 --warmup 2 > $O/r02k2_c4shard.json 2> $O/r02k2_c4shard.err; echo "c4 shard rc=$?"
timeout 600 python bench.py --config c5 --no-cpu --no-extra --steps 6 --warmup 2 > $O/r02k2_c5.json 2> $O/r02k2_c5.err; echo "c5 rc=$?"
python - <<PY
import json
for t in ("c4shard","c5"):
    n=json.load(open("$O/r02k2_%s.json"%t))
    try: o=json.load(open("$O/r02k_%s_old.json"%t))["recall_sweep"]
    except Exception: o={}
    for ef,b in n["recall_sweep"].items():
        a=o.get(ef)
        print("%s ef %5s recall %.4f qps %9.0f gathered %5.0f GB/s"%(t,ef,b["recall_at_k"],b["qps"],b["gathered_gbs_per_gpu"]) + ("  (before: %9.0f, x%.2f)"%(a["qps"], b["qps"]/a["qps"]) if isinstance(a,dict) else ""))
PY
