#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_exact_tc.py tests/test_gpu_exact_stream.py tests/test_gpu_parity.py -x -q -m gpu > $O/r02n_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02n_pytest.log
export CHECK=0 VSANN_BUILD_TC=0
N=1000000 NQ=10000 D=128 K=10 python scripts/tc_probe.py 2>&1 | grep "exact search" | tail -1
N=1000000 NQ=10000 D=128 K=10 SKIP=0 bash scripts/ncu_tc_breakdown.sh > $O/r02n_breakdown.txt 2>&1; head -7 $O/r02n_breakdown.txt
N=400000 NQ=2048 D=768 K=10 METRIC=cos python scripts/tc_probe.py 2>&1 | grep "exact search" | tail -1
N=400000 NQ=2048 D=768 K=10 METRIC=cos SKIP=0 bash scripts/ncu_tc_breakdown.sh 2>&1 | head -7
N=400000 NQ=1024 D=1536 K=100 METRIC=ip python scripts/tc_probe.py 2>&1 | grep "exact search" | tail -1
N=400000 NQ=1024 D=1536 K=100 METRIC=ip SKIP=0 bash scripts/ncu_tc_breakdown.sh 2>&1 | head -7
