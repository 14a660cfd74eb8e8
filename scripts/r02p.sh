#!/bin/bash
# GPU call r02p (2 GPUs): torchrun bench with the ranks' timed steps aligned on the device, and the end-to-end step with the query block
# uploaded in slices + gathered over NVLink against every rank reading the host block in place
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02p_n2.json 2> $O/r02p_n2.err; echo "n2 rc=$?"; tail -3 $O/r02p_n2.err
VSANN_BENCH_ALIGN=0 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02p_n2_noalign.json 2> $O/r02p_n2_noalign.err; echo "n2 noalign rc=$?"
python - <<PY
import json
for t in ("n2","n2_noalign"):
    try:
        b=json.load(open("$O/r02p_%s.json"%t))
        print(t, "value %.2fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
              "e2e", b["e2e"]["path"], b["e2e"]["ms_per_step_by_path"], "recall", b["e2e"]["recall_at_k"])
    except Exception as e:
        print(t, "failed", e)
PY
