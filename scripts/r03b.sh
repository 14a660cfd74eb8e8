#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r03b_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r03b_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 20 --warmup 5 --no-cpu > $O/r03b_bench.json 2> $O/r03b_bench.err; echo "bench rc=$?"
python - <<PY
import json
b=json.load(open("$O/r03b_bench.json"))
print("value %.3fM ms %.4f e2e %.4f build %.3fM/s exact %.2fM qps (%.2f ms, scan %.2f ms frac %.3f) %s"%(b["value"]/1e6,b["ms_per_step"],b["e2e"]["ms_per_step"],b["build"]["vectors_per_s"]/1e6,b["exact"]["qps"]/1e6,b["exact"]["ms_per_batch"],b["exact"]["roofline"]["kernel_ms"],b["exact"]["roofline"]["frac"],b["exact"]["roofline"]["tf32_tflops_measured_here"]))
PY
