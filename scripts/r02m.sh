#!/bin/bash
# GPU call r02m: where does the tensor scan (K1, EMIT pass) spend its time at 1M x 128 x 10k queries? plain timing, then one --set full capture of the
# second k_tc_scan launch of a search (the first is the sampling pass), build batches kept off the tensor path so that the launch index is known.
O=gpurun_out; mkdir -p $O
export N=1000000 NQ=10000 D=128 K=10 CHECK=0 VSANN_BUILD_TC=0
python scripts/tc_probe.py > $O/r02m3_plain.log 2>&1 || { echo plain failed; tail -5 $O/r02m3_plain.log; exit 1; }
cat $O/r02m3_plain.log
ncu --set full --clock-control none --import-source on -k regex:"k_tc_scan" -s 1 -c 1 -f -o $O/r02m3_tcscan python scripts/tc_probe.py > $O/r02m3_ncu.log 2>&1; echo "capture rc=$?"
python scripts/ncu_sum.py $O/r02m3_tcscan.ncu-rep > $O/r02m3_tcscan_fullset_summary.txt 2>&1
cat $O/r02m3_tcscan_fullset_summary.txt | head -40
ncu -i $O/r02m3_tcscan.ncu-rep --page raw --csv > $O/r02m3_tcscan_raw.csv 2>/dev/null
ncu -i $O/r02m3_tcscan.ncu-rep --page source --csv --print-source=cuda,sass > $O/r02m3_tcscan_src.csv 2>/dev/null
python scripts/src_hot.py $O/r02m3_tcscan_src.csv 40 > $O/r02m3_tcscan_source_lines.txt 2>&1
head -45 $O/r02m3_tcscan_source_lines.txt
