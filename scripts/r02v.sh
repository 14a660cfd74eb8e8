#!/bin/bash
# GPU call r02v (4 GPUs): the driver's scaling run visits N = 4 too
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
$TR --master-port 29512 bench.py --gpus 4 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02v_c2_n4.json 2> $O/r02v_c2_n4.err; echo "n4 rc=$?"; grep "search:\|e2e breakdown" $O/r02v_c2_n4.err | cut -c1-400
python - <<PY
import json
b=json.load(open("$O/r02v_c2_n4.json"))
print("value %.2fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
      "global ef", b["operating_point"]["smallest_ef_global"], "e2e", b["e2e"]["ms_per_step"], "build %.2fM/s"%(b["build"]["vectors_per_s"]/1e6))
PY
