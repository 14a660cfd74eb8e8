#!/bin/bash
# GPU call r02g: BASELINE configs[2..4] at FULL size on one B200 (rows generated on the device).
O=gpurun_out; mkdir -p $O
echo "(service tests: covered by the full suite, profiles/r02h)"
timeout 600 python scripts/ingest_bench.py > $O/r02g_ingest.json 2> $O/r02g_ingest.err; echo "ingest rc=$?"; cat $O/r02g_ingest.json | head -c 1500; echo
timeout 900 python bench.py --config c5 --steps 10 --warmup 3 > $O/r02g_c5_n1.json 2> $O/r02g_c5_n1.err; echo "c5 rc=$?"; tail -4 $O/r02g_c5_n1.err
timeout 900 python bench.py --config c4 --capacity-hint 8000000 --steps 10 --warmup 3 > $O/r02g_c4_n1.json 2> $O/r02g_c4_n1.err; echo "c4 rc=$?"; tail -4 $O/r02g_c4_n1.err
timeout 900 python bench.py --config c3 --steps 10 --warmup 3 > $O/r02g_c3_n1.json 2> $O/r02g_c3_n1.err; echo "c3 rc=$?"; tail -4 $O/r02g_c3_n1.err
python - <<PY
import json
for t in ("c5","c4","c3"):
    try:
        b=json.load(open("$O/r02g_%s_n1.json"%t))
        print(t, "value %.3fM"%(b["value"]/1e6), "ms", round(b["ms_per_step"],4), "ef", b["operating_point"]["ef_search"], "recall", b["operating_point"]["recall_at_k"], "frac", round(b["roofline"]["frac"],3),
              "e2e ms", round(b["e2e"]["ms_per_step"],4), "build", round(b["build"]["vectors_per_s"]/1e6,3), "M/s", "growths", b["build"].get("capacity_growths"), "reserve_s", b["build"].get("reserve_seconds"))
    except Exception as e:
        print(t, "failed", e)
PY
