#!/bin/bash
# usage: scripts/ab_env.sh "A=1 B=2" "A=3" ...  -> one bench run per environment setting
for cfg in "$@"; do
  env $cfg python bench.py --steps 20 --warmup 3 --no-cpu 2>gpurun_out/ab.err | tail -1 > gpurun_out/ab.json
  python -c "
import json; r=json.load(open('gpurun_out/ab.json')); c=r['config']; print('== [$cfg]', 'qps', round(r['value']), 'ef', c['ef_search'], 'recall', c['recall_at_10'], 'sweep', {k: c['ef_sweep_recall'][k] for k in ('16','24','32','64')}, 'parity qps', round(r['parity_point']['qps']), 'build', round(r['build']['vectors_per_s']), 'evals', r['roofline']['algorithmic']['dist_evals_per_query'], 'exps', r['roofline']['algorithmic']['expansions_per_query'], 'traffic-frac', r['roofline']['frac'])" || tail -3 gpurun_out/ab.err
done
