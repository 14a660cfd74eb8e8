#!/bin/bash
# GPU call r02c: refactored host layer (shards, VMM tables, concurrent searches, compaction, deterministic inbox) - tests, then bench.
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02c_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r02c_pytest.log
python bench.py --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02c_bench.json 2> $O/r02c_bench.err; echo "bench rc=$?"; tail -3 $O/r02c_bench.err
python - <<PY
import json
b=json.load(open("$O/r02c_bench.json"))
print("ms", round(b["ms_per_step"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"], "frac", round(b["roofline"]["frac"],3))
print("build", json.dumps(b["build"])[:900]); print("e2e", b["e2e"])
PY
