#!/bin/bash
# GPU call r02k: A/B of the wide-beam changes (16-bit-tag visited hash above 1024 slots, launch bounds by list width):
# lib/libvsann.so (before) vs lib/libvsann_w.so (after). Tests first, on the new library.
O=gpurun_out; mkdir -p $O
NEW=$PWD/vector-store-text_b200/lib/libvsann_w.so
VSANN_LIB=$NEW timeout 1200 python -m pytest tests -m gpu -x -q > $O/r02k_pytest_new.log 2>&1; echo "pytest(new lib) rc=$?"; tail -3 $O/r02k_pytest_new.log
for L in old new; do
  if [ $L = new ]; then export VSANN_LIB=$NEW; else unset VSANN_LIB; fi
  VSANN_BENCH_SWEEP=1 timeout 600 python bench.py --config c4 --n 12500000 --no-cpu --no-extra --steps 6 --warmup 2 > $O/r02k_c4shard_$L.json 2> $O/r02k_c4shard_$L.err; echo "c4 shard $L rc=$?"
  timeout 600 python bench.py --config c5 --no-cpu --no-extra --steps 6 --warmup 2 > $O/r02k_c5_$L.json 2> $O/r02k_c5_$L.err; echo "c5 $L rc=$?"
  timeout 300 python bench.py --no-cpu --no-extra --one-build --steps 20 --warmup 5 > $O/r02k_c2_$L.json 2> $O/r02k_c2_$L.err; echo "c2 $L rc=$?"
done
python - <<PY
import json
for t in ("c4shard","c5"):
    o=json.load(open("$O/r02k_%s_old.json"%t)); n=json.load(open("$O/r02k_%s_new.json"%t))
    print(t, "build old %.3f new %.3f M/s"%(o["build"]["vectors_per_s"]/1e6, n["build"]["vectors_per_s"]/1e6))
    for ef in o["recall_sweep"]:
        a,b=o["recall_sweep"][ef], n["recall_sweep"].get(ef)
        if isinstance(a,dict) and isinstance(b,dict):
            print("  ef %5s  recall %.4f -> %.4f   qps %9.0f -> %9.0f (x%.2f)   gathered GB/s %5.0f -> %5.0f   evals/q %6.0f -> %6.0f"%(ef,a["recall_at_k"],b["recall_at_k"],a["qps"],b["qps"],b["qps"]/a["qps"],a["gathered_gbs_per_gpu"],b["gathered_gbs_per_gpu"],a["kernel_dist_evals_per_query"],b["kernel_dist_evals_per_query"]))
o=json.load(open("$O/r02k_c2_old.json")); n=json.load(open("$O/r02k_c2_new.json"))
print("c2 ms old %.4f new %.4f; build old %.3f new %.3f M/s; parity-point qps old %.2fM new %.2fM"%(o["ms_per_step"],n["ms_per_step"],o["build"]["vectors_per_s"]/1e6,n["build"]["vectors_per_s"]/1e6,o["parity_point"]["qps"]/1e6,n["parity_point"]["qps"]/1e6))
PY
