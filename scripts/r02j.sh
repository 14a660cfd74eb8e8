#!/bin/bash
# GPU call r02j: the round's profile set for configs[1] (B200_PROFILING.md). A number printed under ncu is never a bench value:
# 1. plain bench (both arms) -> r02j_bench.json / r02j_reference.json  2. ncu launch list of the same command
# 3. one --set full capture of the search kernel at the ef the bench chose (+ DRAM traffic json)  4. a --set full capture of the tensor scan.
O=gpurun_out; mkdir -p $O
python bench.py --impl reference --steps 20 --warmup 5 > $O/r02j_reference.json 2> $O/r02j_reference.err; echo "reference rc=$?"
python bench.py --steps 20 --warmup 5 > $O/r02j_bench.json 2> $O/r02j_bench.err || { echo "plain bench failed"; tail -5 $O/r02j_bench.err; exit 1; }
EF=$(python -c "import json; print(json.load(open('$O/r02j_bench.json'))['operating_point']['ef_search'])"); echo "bench ok, captures at ef=$EF"
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/r02j_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --one-build --ef $EF > $O/r02j_ncu1.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"^k_search_t" -s 4 -c 1 -f -o $O/r02j_search \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --one-build --ef $EF > $O/r02j_ncu2.log 2>&1; echo "search capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_tc_scan" -s 210 -c 1 -f -o $O/r02j_tcscan \
    python bench.py --steps 2 --warmup 3 --no-cpu --one-build --ef $EF > $O/r02j_ncu3.log 2>&1; echo "tc scan capture rc=$?"
python scripts/ncu_sum.py $O/r02j_search.ncu-rep > $O/r02j_search_fullset_summary.txt 2>&1
python scripts/ncu_sum.py $O/r02j_tcscan.ncu-rep > $O/r02j_tcscan_fullset_summary.txt 2>&1
python scripts/launch_summary.py $O/r02j_launches.csv > $O/r02j_launches_summary.txt 2>&1
ncu -i $O/r02j_search.ncu-rep --page source --csv --print-source=cuda,sass > $O/r02j_search_src.csv 2>/dev/null
python scripts/src_hot.py $O/r02j_search_src.csv 60 > $O/r02j_search_source_lines.txt 2>&1
python - <<PY
import csv, json, subprocess
out = subprocess.run(["ncu", "-i", "$O/r02j_search.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); d = dict(zip(rows[0], rows[2])); u = dict(zip(rows[0], rows[1]))
def b(name):
    v = float(d[name].replace(",", "")); return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u[name]]
r, w = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
json.dump({"kernel": d["Kernel Name"][:80], "ef": $EF, "config": "c2", "dist": "clustered", "dram_bytes_per_launch": r + w, "dram_read": r, "dram_write": w,
           "source": "profiles/r02j_search_fullset_summary.txt (ncu --set full, scripts/r02j.sh: bench.py --steps 2 --warmup 3 --no-cpu --no-extra --one-build --ef $EF, launch #5 of k_search_t)"},
          open("$O/r02_k_search_traffic.json", "w"))
print(open("$O/r02_k_search_traffic.json").read())
PY
head -c 600 $O/r02j_bench.json; echo; cat $O/r02j_launches_summary.txt | head -16
