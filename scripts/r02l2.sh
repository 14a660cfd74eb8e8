#!/bin/bash
# GPU call r02l: insertion batch fraction (VSANN_BATCH_DIV: a batch is at most 1/DIV of the landed graph) against build rate and graph quality,
# and the current phase split of the build (VSANN_BUILD_TRACE).
O=gpurun_out; mkdir -p $O
for DIV in 6 4 2; do
  VSANN_BATCH_DIV=$DIV timeout 300 python bench.py --no-cpu --no-extra --one-build --steps 10 --warmup 3 > $O/r02l_div$DIV.json 2> $O/r02l_div$DIV.err; echo "div $DIV rc=$?"
done
python - <<PY
import json
for d in (6,4,2):
    b=json.load(open("$O/r02l_div%d.json"%d)); sw=b["operating_point"]["ef_sweep_recall"]
    print("div",d,"build %.3f M/s wall (%d batches)"%(b["build"]["vectors_per_s"]/1e6,b["build"]["batches"]),"ef",b["operating_point"]["ef_search"],"ms %.4f"%b["ms_per_step"],
          "recall@ef16/18/20/24/64:",sw.get("16"),sw.get("18"),sw.get("20"),sw.get("24"),sw.get("64"),"evals/q %.0f"%b["roofline"]["algorithmic"]["dist_evals_per_query"], "parity qps %.2fM"%(b["parity_point"]["qps"]/1e6))
PY
