#!/bin/bash
# GPU call r02i (8 GPUs of one box): BASELINE configs[2] and configs[3] at FULL size over 8 shards, configs[1] x 8 (weak scaling,
# fused-flag exchange) under torchrun and as ONE 8-device vs_index in one process, and the peer check on 8 real peers.
O=gpurun_out; mkdir -p $O
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29521 scripts/peer_check.py > $O/r02i_peer_check_n$N.log 2>&1; echo "peer_check rc=$?"; grep -E "PEER CHECK" $O/r02i_peer_check_n$N.log
timeout 900 $TR --master-port 29522 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02i_c2_n$N.json 2> $O/r02i_c2_n$N.err; echo "c2 torchrun rc=$?"
timeout 900 python bench.py --gpus $N --single-process --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02i_c2_n${N}_single.json 2> $O/r02i_c2_n${N}_single.err; echo "c2 single-process rc=$?"; tail -2 $O/r02i_c2_n${N}_single.err
timeout 900 $TR --master-port 29523 bench.py --gpus $N --config c3 --steps 10 --warmup 3 --no-cpu > $O/r02i_c3_n$N.json 2> $O/r02i_c3_n$N.err; echo "c3 rc=$?"; tail -2 $O/r02i_c3_n$N.err
timeout 900 $TR --master-port 29524 bench.py --gpus $N --config c4 --capacity-hint 2000000 --steps 10 --warmup 3 --no-cpu > $O/r02i_c4_n$N.json 2> $O/r02i_c4_n$N.err; echo "c4 rc=$?"; tail -2 $O/r02i_c4_n$N.err
python - <<PY
import json
for t in ("c2_n$N","c2_n${N}_single","c3_n$N","c4_n$N"):
    try:
        b=json.load(open("$O/r02i_%s.json"%t))
        print(t, "value %.3fM q/s"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
              "frac", round(b["roofline"]["frac"],3), "e2e ms", round(b["e2e"]["ms_per_step"],4), "build %.3f M/s"%(b["build"]["vectors_per_s"]/1e6), "growths", b["build"].get("capacity_growths"), "reserve_s", b["build"].get("reserve_seconds"))
    except Exception as e:
        print(t, "failed", e)
PY
