#!/bin/bash
# usage: env <probe vars> scripts/ncu_tc_breakdown.sh  -> per-kernel time of one exact-search call through the tcgen05 path
ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"k_tc_|k_query_norms_tc|k_exact_tiles<[0-9], .*LessByKey" -s ${SKIP:-0} -c 16 --csv --log-file gpurun_out/tcb.csv python scripts/tc_probe.py > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/tcb.csv")) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); mi=h.index("Metric Name")
out={}
for r in rows[1:]:
    out.setdefault((r[h.index("ID")], r[ki][:44]), {})[r[mi][:12]] = r[vi]
for (i,k),v in out.items(): print(f"{k:46s} {float(v.get('gpu__time_du',0))/1e6:9.3f} ms  tensor {v.get('sm__pipe_ten','-')}%")
PY
