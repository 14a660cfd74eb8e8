#!/bin/bash
# GPU call r03d (8 GPUs): configs[1] x 8 under torchrun with the round's final library, full default line (CPU leg on rank 0 skipped to save box time)
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu > $O/r03d_c2_n8.json 2> $O/r03d_c2_n8.err; echo "n8 rc=$?"; grep "search:" $O/r03d_c2_n8.err
python - <<PY
import json
b=json.load(open("$O/r03d_c2_n8.json"))
print("value %.2fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
      "global ef", b["operating_point"]["smallest_ef_global"], "e2e", round(b["e2e"]["ms_per_step"],4), b["e2e"]["per_step_spread"], "build %.2fM/s"%(b["build"]["vectors_per_s"]/1e6), "exact qps", b["exact"]["qps"] if b.get("exact") else None)
PY
