#!/bin/bash
# ncu --set full of one launch of a kernel of a workload: tools/gpu_ncu_wl.sh <workload> <kernel-regex> <tag> [skip]
WL=$1; RX=$2; TAG=$3; SKIP=${4:-6}
CMD="python bench.py --workload $WL --steps 8 --warmup 3 --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/ncu_$TAG.log
