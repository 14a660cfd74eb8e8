#!/bin/bash
# GPU call r02o: whole GPU suite + the default bench after the tensor-scan rework
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r02o_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02o_pytest.log
python bench.py --steps 20 --warmup 5 > $O/r02o_bench.json 2> $O/r02o_bench.err; echo "bench rc=$?"; tail -3 $O/r02o_bench.err
python - <<PY
import json
b=json.load(open("$O/r02o_bench.json"))
print("value %.3f M qps, ms %.4f, e2e %.4f ms, build %.3f M/s wall"%(b["value"]/1e6,b["ms_per_step"],b["e2e"]["ms_per_step"],b["build"]["vectors_per_s"]/1e6))
print("exact", {k:v for k,v in b["exact"].items() if k in ("qps","ms_per_batch","second_scan_queries","fp32_fallback_queries")}, b["exact"]["roofline"])
print("iid", b.get("iid_gaussian",{}).get("build"), )
PY
