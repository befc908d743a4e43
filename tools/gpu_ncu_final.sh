#!/bin/bash
# final captures of the two kernels of the headline workload, a mid-path launch each (ncu --set full, after the same
# command has run without ncu)
CMD="python bench.py --workload config4_aopp_4k --steps 40 --warmup 8 --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check"
$CMD > gpurun_out/plain_final.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:rt_ao_postprocess_kernel -s 30 -c 1 -f -o gpurun_out/prof_pp_final $CMD > gpurun_out/ncu_pp_final.log 2>&1; echo "pp rc=$?"
ncu --set full --clock-control none --import-source on -k regex:rt_ao_pulled_kernel -s 30 -c 1 -f -o gpurun_out/prof_pulled_final $CMD > gpurun_out/ncu_pulled_final.log 2>&1; echo "pulled rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_final_config4.csv python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check > gpurun_out/ncu_ll_final.log 2>&1; echo "launch list rc=$?"
