#!/bin/bash
# GPU call r02s: the round's final profile set for configs[1] (B200_PROFILING.md). A number printed under ncu is never a bench value:
# 1. whole GPU suite + smoke()  2. plain bench, both arms  3. ncu launch list of the same command
# 4. one --set full capture of the search kernel at the ef the bench chose (+ DRAM traffic json)  5. a --set full capture of the tensor scan (EMIT pass).
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r02s_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02s_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py --impl reference --steps 20 --warmup 5 > $O/r02s_reference.json 2> $O/r02s_reference.err; echo "reference rc=$?"
python bench.py --steps 20 --warmup 5 > $O/r02s_bench.json 2> $O/r02s_bench.err || { echo "plain bench failed"; tail -5 $O/r02s_bench.err; exit 1; }
EF=$(python -c "import json; print(json.load(open('$O/r02s_bench.json'))['operating_point']['ef_search'])"); echo "bench ok, captures at ef=$EF"
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/r02s_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --one-build --ef $EF > $O/r02s_ncu1.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"^k_search_t" -s 4 -c 1 -f -o $O/r02s_search \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --one-build --ef $EF > $O/r02s_ncu2.log 2>&1; echo "search capture rc=$?"
N=1000000 NQ=10000 D=128 K=10 CHECK=0 VSANN_BUILD_TC=0 ncu --set full --clock-control none --import-source on -k regex:"k_tc_scan" -s 1 -c 1 -f -o $O/r02s_tcscan python scripts/tc_probe.py > $O/r02s_ncu3.log 2>&1; echo "tc scan capture rc=$?"
python scripts/ncu_sum.py $O/r02s_search.ncu-rep > $O/r02s_search_fullset_summary.txt 2>&1
python scripts/ncu_sum.py $O/r02s_tcscan.ncu-rep > $O/r02s_tcscan_fullset_summary.txt 2>&1
python scripts/launch_summary.py $O/r02s_launches.csv > $O/r02s_launches_summary.txt 2>&1
ncu -i $O/r02s_search.ncu-rep --page source --csv --print-source=cuda,sass > $O/r02s_search_src.csv 2>/dev/null
python scripts/src_hot.py $O/r02s_search_src.csv 60 > $O/r02s_search_source_lines.txt 2>&1
ncu -i $O/r02s_tcscan.ncu-rep --page source --csv --print-source=cuda,sass > $O/r02s_tcscan_src.csv 2>/dev/null
python scripts/src_hot.py $O/r02s_tcscan_src.csv 40 > $O/r02s_tcscan_source_lines.txt 2>&1
python - <<PY
import csv, json, subprocess
out = subprocess.run(["ncu", "-i", "$O/r02s_search.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); d = dict(zip(rows[0], rows[2])); u = dict(zip(rows[0], rows[1]))
def b(name):
    v = float(d[name].replace(",", "")); return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u[name]]
r, w = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
json.dump({"kernel": d["Kernel Name"][:80], "ef": $EF, "config": "c2", "dist": "clustered", "dram_bytes_per_launch": r + w, "dram_read": r, "dram_write": w,
           "source": "profiles/r02s_search_fullset_summary.txt (ncu --set full, scripts/r02s.sh: bench.py --steps 2 --warmup 3 --no-cpu --no-extra --one-build --ef $EF, launch #5 of k_search_t)"},
          open("$O/r02_k_search_traffic.json", "w"))
print(open("$O/r02_k_search_traffic.json").read())
PY
head -c 400 $O/r02s_bench.json; echo; head -14 $O/r02s_launches_summary.txt
