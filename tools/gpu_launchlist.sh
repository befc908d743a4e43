#!/bin/bash
# per-launch durations of a short bench run: tools/gpu_launchlist.sh <kernel env> [bench args]
k=$1; shift
CMD="python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu-baseline $@"
RT_AO_KERNEL=$k $CMD > gpurun_out/plain_ll.log 2>&1 &&
RT_AO_KERNEL=$k ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__icc_request_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 12 -c 9 --csv --log-file gpurun_out/ll_$k.csv $CMD > gpurun_out/ncu_ll.log 2>&1
python - <<PY
import csv
rows=list(csv.DictReader(l for l in open("gpurun_out/ll_$k.csv") if l.startswith('"')))
agg={}
for r in rows:
    agg.setdefault((r["ID"], r["Kernel Name"].split("(")[0]), {})[r["Metric Name"]]=r["Metric Value"]
for (i,n),m in agg.items():
    print(i, n, {k.split("__")[-1][:28]:v for k,v in m.items()})
PY
