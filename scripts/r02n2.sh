#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_exact_tc.py tests/test_gpu_exact_stream.py tests/test_gpu_parity.py -x -q -m gpu > $O/r02n_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02n_pytest.log
N=100000 NQ=1000 D=200 K=10 CHECK=1 METRIC=cos VSANN_EXACT_TC=1 python scripts/tc_probe.py 2>&1 | tail -1
export N=1000000 NQ=10000 D=128 K=10 CHECK=0 VSANN_BUILD_TC=0
for F in 0 1 2 4; do echo "flags $F"; VSANN_TC_FLAGS=$F python scripts/tc_probe.py 2>&1 | grep "exact search" | tail -1; done
SKIP=0 bash scripts/ncu_tc_breakdown.sh > $O/r02n_breakdown.txt 2>&1; head -7 $O/r02n_breakdown.txt
