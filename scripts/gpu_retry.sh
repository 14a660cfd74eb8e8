#!/bin/bash
# usage: scripts/gpu_retry.sh <gpus> <timeout> <script>   - retries while the pod answers busy (nothing is charged for those)
G=$1; T=$2; S=$3
for i in $(seq 1 30); do
  if [ "$G" = "1" ]; then OUT=$(gpurun --timeout $T -- "bash $S" 2>&1); else OUT=$(gpurun --gpus $G --timeout $T -- "bash $S" 2>&1); fi
  if echo "$OUT" | grep -q "status=transient"; then echo "[retry $i] busy"; sleep 150; continue; fi
  echo "$OUT" | tail -60; exit 0
done
echo "gave up"
