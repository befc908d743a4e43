#!/bin/bash
# Round-2 profile set (one GPU): the launch list of the bench command, then one `ncu --set full` capture of each of
# the four final kernels on the workload that exercises it.  Each capture runs only after the same command has
# exited 0 without ncu (tools/gpu_ncu_wl.sh).
CMD="python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check"
$CMD > gpurun_out/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_config4.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
bash tools/gpu_ncu_wl.sh config4_aopp_4k rt_ao_pulled_kernel pulled_config4 8
bash tools/gpu_ncu_wl.sh config4_aopp_4k rt_ao_postprocess_kernel pp_config4 8
bash tools/gpu_ncu_wl.sh config2_ao_1080p rt_ao_pulled_kernel pulled_config2 6
bash tools/gpu_ncu_wl.sh config3_hybrid_1080p rt_hybrid_kernel hybrid_config3 6
bash tools/gpu_ncu_wl.sh config1_phong_640x480 rt_phong_kernel phong_config1 6
ls -la gpurun_out/*.ncu-rep
