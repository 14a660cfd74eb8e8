#!/bin/bash
# GPU call r02e (2 GPUs): peers check, multi-device index test, torchrun bench (flag barrier vs torch barrier), single-process bench.
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29511 scripts/peer_check.py > $O/r02e_peer_check.log 2>&1; echo "peer_check rc=$?"; grep -E "PEER CHECK|flags=" $O/r02e_peer_check.log | tail -8
timeout 600 python -m pytest tests/test_gpu_mutations.py -m gpu -q -k "one_index_over_devices" -s > $O/r02e_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -4 $O/r02e_pytest_multi.log
$TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02e_bench_n2_flags.json 2> $O/r02e_bench_n2_flags.err; echo "n2 flags rc=$?"
VSANN_P2P_FLAGS=0 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02e_bench_n2_barrier.json 2> $O/r02e_bench_n2_barrier.err; echo "n2 barrier rc=$?"
python bench.py --gpus 2 --single-process --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02e_bench_n2_single.json 2> $O/r02e_bench_n2_single.err; echo "n2 single rc=$?"; tail -3 $O/r02e_bench_n2_single.err
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu --no-extra > $O/r02e_bench_n1.json 2> $O/r02e_bench_n1.err; echo "n1 rc=$?"
python - <<PY
import json
for t in ("n1","n2_flags","n2_barrier","n2_single"):
    try:
        b=json.load(open("$O/r02e_bench_%s.json"%t))
        print(t, "value %.2fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "kernel", round(b["roofline"]["kernel_ms"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"],
              "e2e ms", round(b["e2e"]["ms_per_step"],4), "build", round(b["build"]["vectors_per_s"]/1e6,3), "M/s")
    except Exception as e:
        print(t, "failed", e)
PY
