#!/bin/bash
# one `--set full` capture of the timed search kernel (run only after the same command exited 0 without ncu)
TAG=${1:-cap}; O=gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --ef 24 > $O/${TAG}_plain.log 2>&1 || { echo plain failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"^k_search_t" -s 4 -c 1 -f -o $O/${TAG}_search \
    python bench.py --steps 2 --warmup 3 --no-cpu --ef 24 > $O/${TAG}_ncu.log 2>&1; echo "capture rc=$?"
ncu -i $O/${TAG}_search.ncu-rep --page source --print-source=sass --csv > $O/${TAG}_search_sass.csv 2>/dev/null
ncu -i $O/${TAG}_search.ncu-rep --page source --print-source=cuda,sass --csv > $O/${TAG}_search_src.csv 2>/dev/null
ls -la $O | grep ${TAG}
