#!/bin/bash
# GPU call r02h: the allocation-free entry point alone in a fresh process, the ingest test repeated, then the whole GPU suite with -x.
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/r02h_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -4 $O/r02h_noalloc_alone.log
for i in 1 2 3 4 5; do timeout 300 python -m pytest tests/test_gpu_service.py -m gpu -q -k out_of_order > $O/r02h_ingest_$i.log 2>&1; echo "ingest run $i rc=$?"; done
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02h_pytest.log 2>&1; echo "full suite (-x) rc=$?"; tail -8 $O/r02h_pytest.log
