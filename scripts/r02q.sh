#!/bin/bash
# GPU call r02q (8 GPUs): configs[1] x 8 under torchrun after the rank merge / parallel flag polling; both end-to-end paths
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29511 scripts/peer_check.py > $O/r02q_peer_check_n8.log 2>&1; echo "peer_check rc=$?"; grep -E "PEER CHECK|exchange step alone|sharded step" $O/r02q_peer_check_n8.log | tail -8
$TR --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02q_c2_n8.json 2> $O/r02q_c2_n8.err; echo "n8 rc=$?"; tail -2 $O/r02q_c2_n8.err
python - <<PY
import json
b=json.load(open("$O/r02q_c2_n8.json"))
print("value %.2fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
      "global ef", b["operating_point"]["smallest_ef_global"], "e2e", b["e2e"]["ms_per_step_by_path"], b["e2e"]["breakdown"])
PY
