#!/bin/bash
# GPU call r02b: A/B of the unsorted register list (REDUX) against the sorted list; correctness first.
O=gpurun_out; mkdir -p $O
U=$PWD/vector-store-text_b200/lib/libvsann_u.so
VSANN_LIB=$U timeout 600 python -m pytest tests -m gpu -x -q > $O/r02b_pytest_u.log 2>&1; echo "pytest(new list) rc=$?"; tail -3 $O/r02b_pytest_u.log
python bench.py --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02b_bench_old.json 2> $O/r02b_bench_old.err; echo "old rc=$?"
VSANN_LIB=$U python bench.py --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02b_bench_new.json 2> $O/r02b_bench_new.err; echo "new rc=$?"
python - <<PY
import json
for t in ("old","new"):
    b=json.load(open("$O/r02b_bench_%s.json"%t))
    print(t, "ms", round(b["ms_per_step"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"], "frac", round(b["roofline"]["frac"],3),
          "parity qps", round(b["parity_point"]["qps"]/1e6,2), "build dev", b["build"]["device_resident"]["vectors_per_s"], "build wall", b["build"]["vectors_per_s"], "e2e ms", b["e2e"]["ms_per_step"])
PY
