#!/bin/bash
# phased AO kernel block-size variants at config 2 / config 4: rebuild with -D flags on the GPU box
for v in "-DPH_THREADS_N=128" "-DPH_THREADS_N=64" "-DPH_THREADS_N=32" "-DPH_THREADS_N=256"; do
  RT_NVCC_EXTRA="$v" python -m real_time_ray_tracer_b200.build --force > /dev/null 2>&1 || { echo "build failed $v"; continue; }
  for wl in config2_ao_1080p config5_aopp_8k_aa16; do
  st=256; [ $wl = config5_aopp_8k_aa16 ] && st=8
  python bench.py --workload $wl --steps $st --warmup 8 --no-e2e --no-cpu-baseline > gpurun_out/var.json 2>gpurun_out/var.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/var.json").read().strip().splitlines()[-1])
print("$v $wl", "frame", d["ms_per_step"], "trace_ms", d["roofline"]["kernel_ms"])
PY
  done
done
python -m real_time_ray_tracer_b200.build --force > /dev/null 2>&1
