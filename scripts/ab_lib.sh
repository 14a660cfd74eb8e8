#!/bin/bash
# A/B of alternative builds of the library (lib/alt/libvsann_<tag>.so) on the bench workload
for tag in ${TAGS:-default}; do
  if [ "$tag" = default ]; then unset VSANN_LIB; else export VSANN_LIB=$PWD/vector-store-text_b200/lib/alt/libvsann_$tag.so; fi
  python bench.py --steps 20 --warmup 3 --no-cpu --ef 24 2>gpurun_out/ab.err | tail -1 > gpurun_out/ab.json
  python -c "
import json; r=json.load(open('gpurun_out/ab.json')); print('== lib $tag', r['value'], r['ms_per_step'], r['config'].get('recall_at_10'), r['e2e']['value'], r['build']['vectors_per_s'])" || tail -3 gpurun_out/ab.err
done
