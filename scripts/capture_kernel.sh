#!/bin/bash
# usage: scripts/capture_kernel.sh TAG KERNEL_REGEX SKIP   -> one `--set full` capture (+ per-source-line csv) of a build kernel
TAG=$1; RE=$2; SKIP=${3:-100}; O=gpurun_out
R=1 python scripts/build_probe.py > $O/${TAG}_plain.log 2>&1 || { echo plain failed; tail -3 $O/${TAG}_plain.log; exit 1; }
R=1 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c 1 -f -o $O/${TAG} python scripts/build_probe.py > $O/${TAG}_ncu.log 2>&1; echo "capture rc=$?"
ncu -i $O/${TAG}.ncu-rep --page source --print-source=cuda,sass --csv > $O/${TAG}_src.csv 2>/dev/null
ls -la $O | grep ${TAG}
