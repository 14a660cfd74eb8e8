#!/bin/bash
# GPU call r02a: new same-config bench (both arms), build trace, late-batch build captures.
O=gpurun_out; mkdir -p $O
nproc > $O/r02a_host.txt; lscpu | grep "Model name" >> $O/r02a_host.txt; nvidia-smi -L >> $O/r02a_host.txt
python bench.py --steps 20 --warmup 5 > $O/r02a_bench.json 2> $O/r02a_bench.err; echo "bench rc=$?"; tail -3 $O/r02a_bench.err
python bench.py --impl reference --steps 20 --warmup 5 > $O/r02a_reference.json 2> $O/r02a_reference.err; echo "reference rc=$?"
VSANN_BUILD_TRACE=1 python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/r02a_trace.json 2> $O/r02a_trace.err; echo "trace rc=$?"; grep "build trace" $O/r02a_trace.err | tail -4
ncu --set full --clock-control none --import-source on -k regex:"^(k_build_search_t|k_select|k_reverse_merge)" -s 600 -c 3 -f -o $O/r02a_build_late \
    python bench.py --steps 1 --warmup 1 --no-cpu --no-extra > $O/r02a_ncu_build.log 2>&1; echo "build capture rc=$?"
for i in 0 1 2; do python scripts/ncu_sum.py $O/r02a_build_late.ncu-rep $i >> $O/r02a_build_late_summary.txt 2>&1; echo >> $O/r02a_build_late_summary.txt; done
head -c 1500 $O/r02a_bench.json; echo
