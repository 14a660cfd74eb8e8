#!/bin/bash
# GPU call r03a: tensor scan as 2-CTA clusters with the corpus tile multicast (VSANN_TC_CLUSTER=0: off): correctness, then timing by shape
O=gpurun_out; mkdir -p $O
N=100000 NQ=2048 D=128 K=10 CHECK=1 VSANN_EXACT_TC=1 timeout 120 python scripts/tc_probe.py 2>&1 | tail -2; echo "probe rc=$?"
N=100000 NQ=1100 D=200 K=10 CHECK=1 METRIC=cos VSANN_EXACT_TC=1 timeout 120 python scripts/tc_probe.py 2>&1 | tail -2; echo "probe rc=$?"
timeout 600 python -m pytest tests/test_gpu_exact_tc.py tests/test_gpu_exact_stream.py tests/test_gpu_parity.py -x -q -m gpu > $O/r03a_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r03a_pytest.log
export CHECK=0 VSANN_BUILD_TC=0
for CL in 1 0; do
echo "cluster=$CL"
VSANN_TC_CLUSTER=$CL N=1000000 NQ=10000 D=128 K=10 SKIP=0 timeout 200 bash scripts/ncu_tc_breakdown.sh 2>&1 | grep "k_tc_scan<1" | head -1
VSANN_TC_CLUSTER=$CL N=400000 NQ=2048 D=768 K=10 METRIC=cos SKIP=0 timeout 200 bash scripts/ncu_tc_breakdown.sh 2>&1 | grep "k_tc_scan<1" | head -1
VSANN_TC_CLUSTER=$CL N=400000 NQ=1024 D=1536 K=100 METRIC=ip SKIP=0 timeout 200 bash scripts/ncu_tc_breakdown.sh 2>&1 | grep "k_tc_scan<1" | head -1
done
