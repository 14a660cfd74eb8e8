#!/bin/bash
# what the driver runs at the end of a round, on one GPU: the GPU suite, smoke(), a short bench
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 5 --warmup 3 --no-cpu --no-extra > $O/final_bench.json 2> $O/final_bench.err; echo "bench rc=$?"; grep "search:" $O/final_bench.err
