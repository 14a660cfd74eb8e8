#!/bin/bash
O=gpurun_out; mkdir -p $O
VSANN_BENCH_SWEEP=1 python bench.py --steps 12 --warmup 3 --no-cpu --no-extra --one-build > $O/r02r_c2_sweep.json 2> $O/r02r_c2_sweep.err; echo "rc=$?"
python - <<PY
import json
b=json.load(open("$O/r02r_c2_sweep.json"))
for ef,v in b["recall_sweep"].items():
    if isinstance(v,dict): print("ef",ef,"recall",v["recall_at_k"],"ms %.4f"%v["ms_per_step"],"qps %.2fM"%(v["qps"]/1e6),"evals/q %.0f"%v["kernel_dist_evals_per_query"],"gathered",round(v["gathered_gbs_per_gpu"]))
PY
