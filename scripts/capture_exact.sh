#!/bin/bash
# usage (GPU box): scripts/capture_exact.sh TAG  -> `--set full` captures of the streaming exact scan (B=1) and the tensor scan (B=64) at 2M x 1536
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
ARGS="--config c5 --only-exact --M 2 --ef-add 8 --n 2000000 --reps 1"
python scripts/config_sweep.py $ARGS --nq 1 --batches 1 > $O/${TAG}_plain1.json 2> $O/${TAG}_plain1.err || { echo plain failed; tail -3 $O/${TAG}_plain1.err; exit 1; }
grep "^# exact" $O/${TAG}_plain1.err
ncu --set full --clock-control none --import-source on -k regex:"k_exact_stream" -s 1 -c 1 -f -o $O/${TAG}_stream python scripts/config_sweep.py $ARGS --nq 1 --batches 1 > $O/${TAG}_ncu1.log 2>&1; echo "stream capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_tc_scan" -s 3 -c 1 -f -o $O/${TAG}_tcscan python scripts/config_sweep.py $ARGS --nq 64 --batches 64 > $O/${TAG}_ncu2.log 2>&1; echo "tc capture rc=$?"
for k in stream tcscan; do python scripts/ncu_sum.py $O/${TAG}_$k.ncu-rep > $O/${TAG}_${k}_fullset_summary.txt 2>&1; done
ls -la $O | grep ${TAG}
