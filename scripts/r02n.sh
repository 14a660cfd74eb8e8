#!/bin/bash
# GPU call r02n: tensor scan after the resident query tile / coefficient ring / deferred candidate writes: exact-path tests, then timing
# at 1M x 128 x 10k (+ oracle check at 100k), then the per-kernel breakdown.
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_exact_tc.py tests/test_gpu_exact_stream.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu > $O/r02n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02n_pytest.log
N=100000 NQ=2048 D=128 K=10 CHECK=1 VSANN_EXACT_TC=1 python scripts/tc_probe.py 2>&1 | tail -3
N=100000 NQ=1000 D=200 K=10 CHECK=1 METRIC=cos VSANN_EXACT_TC=1 python scripts/tc_probe.py 2>&1 | tail -2
export N=1000000 NQ=10000 D=128 K=10 CHECK=0 VSANN_BUILD_TC=0
python scripts/tc_probe.py > $O/r02n_plain.log 2>&1; cat $O/r02n_plain.log
SKIP=0 bash scripts/ncu_tc_breakdown.sh > $O/r02n_breakdown.txt 2>&1; head -12 $O/r02n_breakdown.txt
