#!/bin/bash
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29511 scripts/peer_check.py > $O/r02f_peer_check.log 2>&1; echo "peer_check rc=$?"; grep -E "PEER CHECK|flags=" $O/r02f_peer_check.log | tail -8
$TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02f_bench_n2_flags.json 2> $O/r02f_bench_n2_flags.err; echo "n2 flags rc=$?"
VSANN_P2P_FLAGS=0 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02f_bench_n2_barrier.json 2> $O/r02f_bench_n2_barrier.err; echo "n2 barrier rc=$?"
python - <<PY
import json
for t in ("n2_flags","n2_barrier"):
    try:
        b=json.load(open("$O/r02f_bench_%s.json"%t))
        print(t, "value %.2fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
              "e2e ms", round(b["e2e"]["ms_per_step"],4), "build", round(b["build"]["vectors_per_s"]/1e6,3), "M/s")
    except Exception as e:
        print(t, "failed", e)
PY
