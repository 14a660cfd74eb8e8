#!/bin/bash
# usage (on the GPU box): scripts/profile_round.sh TAG
# 1. bench.py without a profiler -> gpurun_out/TAG_bench.json   (the only numbers that count)
# 2. ncu launch list of the same command (B200_PROFILING.md: gpu__time_duration.sum, no clock control)
# 3. one `--set full` capture of the search traversal kernel and one of the build traversal kernel, with SASS.
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err || { echo "plain bench failed"; tail -5 $O/${TAG}_bench.err; exit 1; }
tail -c 600 $O/${TAG}_bench.json; echo
EF=$(python -c "import json; print(json.load(open('$O/${TAG}_bench.json'))['config']['ef_search'])")   # the captures use the ef the bench settled on
echo "captures at ef=$EF"
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > $O/${TAG}_ncu1.log 2>&1; echo "launch list rc=$?"
# the search kernel: skip the build-time launches by name, take one launch after the sweep's warm-ups
ncu --set full --clock-control none --import-source on -k regex:"^k_search_t" -s 4 -c 1 -f -o $O/${TAG}_search \
    python bench.py --steps 2 --warmup 3 --no-cpu --ef $EF > $O/${TAG}_ncu2.log 2>&1; echo "search capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"^k_build_search_t" -s 40 -c 1 -f -o $O/${TAG}_build \
    python bench.py --steps 1 --warmup 3 --no-cpu --ef $EF > $O/${TAG}_ncu3.log 2>&1; echo "build capture rc=$?"
ncu -i $O/${TAG}_search.ncu-rep --page source --print-source=sass --csv > $O/${TAG}_search_sass.csv 2>/dev/null
python scripts/ncu_sum.py $O/${TAG}_search.ncu-rep > $O/${TAG}_search_fullset_summary.txt
python scripts/ncu_sum.py $O/${TAG}_build.ncu-rep > $O/${TAG}_build_fullset_summary.txt
python scripts/launch_summary.py $O/${TAG}_launches.csv > $O/${TAG}_launches_summary.txt
python - <<EOF
import csv, json, subprocess
out = subprocess.run(["ncu", "-i", "$O/${TAG}_search.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); d = dict(zip(rows[0], rows[2])); u = dict(zip(rows[0], rows[1]))
def b(name):
    v = float(d[name].replace(",", "")); return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u[name]]
r, w = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
json.dump({"kernel": d["Kernel Name"][:80], "ef": $EF, "dram_bytes_per_launch": r + w, "dram_read": r, "dram_write": w,
           "source": "$O/${TAG}_search.ncu-rep (ncu --set full, scripts/profile_round.sh: bench.py --steps 2 --warmup 3 --no-cpu --ef $EF, launch #5 of k_search_t)"},
          open("$O/${TAG}_k_search_traffic.json", "w"))
print(open("$O/${TAG}_k_search_traffic.json").read())
EOF
ls -la $O | grep ${TAG}
