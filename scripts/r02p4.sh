#!/bin/bash
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02p_n2.json 2> $O/r02p_n2.err; echo "n2 rc=$?"; grep "search:\|e2e breakdown" $O/r02p_n2.err
VSANN_P2P_PUSH=0 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02p_n2_pull.json 2> $O/r02p_n2_pull.err; echo "n2 pull rc=$?"; grep "search:\|e2e breakdown" $O/r02p_n2_pull.err
python - <<PY
import json
for t in ("n2","n2_pull"):
    b=json.load(open("$O/r02p_%s.json"%t))
    print(t, "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "e2e", b["e2e"]["ms_per_step_by_path"])
PY
