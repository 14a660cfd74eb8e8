#!/bin/bash
# A/B of the per-warp staging size on the bench workload (build once per setting: the graph depends only on the data)
for sb in ${STAGES:-4096 6144}; do
  VSANN_STAGE_BYTES=$sb python bench.py --steps 20 --warmup 5 --no-cpu 2>/dev/null | tail -1 > gpurun_out/ab.json
  python -c "
import json; r=json.load(open('gpurun_out/ab.json')); print('== stage $sb', r['value'], r['ms_per_step'], r['config'].get('recall_at_10'), r['e2e']['value'], r['config'].get('build_vectors_per_s'))"
done
