#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/r02d_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/r02d_pytest.log
