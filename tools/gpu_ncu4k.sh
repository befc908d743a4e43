#!/bin/bash
# ncu --set full of one AO trace launch and one post-processing launch of config 4 (3840x2160, mode 1)
CMD="python bench.py --workload config4_aopp_4k --steps 8 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"rt_ao_(pulled|phased|postprocess)" -s 16 -c 2 -f -o gpurun_out/prof_c4 $CMD > gpurun_out/ncu_c4.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_c4.log
# launch list of the default bench command (cold-cache, serialised: shares only)
CMD2="python bench.py --steps 16 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD2 > gpurun_out/plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 3 -c 40 --csv --log-file gpurun_out/launches_c2.csv $CMD2 > gpurun_out/ncu_l2.log 2>&1
$CMD > gpurun_out/plain_c4b.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 9 -c 60 --csv --log-file gpurun_out/launches_c4.csv $CMD > gpurun_out/ncu_l4.log 2>&1
echo done
