#!/bin/bash
# usage: BENCH_ARGS="--nq 40000 --ef 24" scripts/ab_env2.sh "A=1" "B=2"
for cfg in "$@"; do
  env $cfg python bench.py --steps 10 --warmup 3 --no-cpu $BENCH_ARGS 2>gpurun_out/ab.err | tail -1 > gpurun_out/ab.json
  python -c "
import json; r=json.load(open('gpurun_out/ab.json')); c=r['config']; print('== [$cfg] $BENCH_ARGS', 'qps', round(r['value']), 'ms', round(r['ms_per_step'],4), 'recall', c['recall_at_10'], 'evals', r['roofline']['algorithmic']['dist_evals_per_query'])" || tail -3 gpurun_out/ab.err
done
