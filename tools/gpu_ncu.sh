#!/bin/bash
# ncu --set full of one launch of the AO kernel selected by $1 (pulled|phased|queued|tile|stream); report -> gpurun_out/prof_$1.ncu-rep
k=${1:-pulled}; shift
CMD="python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu-baseline ${@}"
RT_AO_KERNEL=$k $CMD > gpurun_out/plain_$k.log 2>&1 &&
RT_AO_KERNEL=$k ncu --set full --clock-control none --import-source on -k regex:rt_ao -s 5 -c 1 -f -o gpurun_out/prof_$k $CMD > gpurun_out/ncu_$k.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_$k.log
