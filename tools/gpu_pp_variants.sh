#!/bin/bash
# post-processing kernel variants at config 4: rebuild with -D flags on the GPU box, print kernel times
for v in "-DPP_SPECULATE=1 -DPP_MIN_BLOCKS=10 -DPP_BATCH=2" "-DPP_SPECULATE=1 -DPP_MIN_BLOCKS=12 -DPP_BATCH=2" "-DPP_SPECULATE=1 -DPP_MIN_BLOCKS=12 -DPP_BATCH=1" "-DPP_SPECULATE=1 -DPP_MIN_BLOCKS=10 -DPP_BATCH=1" "-DPP_SPECULATE=0 -DPP_MIN_BLOCKS=10 -DPP_BATCH=2" "-DPP_SPECULATE=1 -DPP_MIN_BLOCKS=9 -DPP_BATCH=2"; do
  RT_NVCC_EXTRA="$v" python -m real_time_ray_tracer_b200.build --force > /dev/null 2>&1 || { echo "build failed $v"; continue; }
  python bench.py --workload config4_aopp_4k --steps 64 --warmup 16 --no-e2e --no-cpu-baseline > gpurun_out/pp_var.json 2>gpurun_out/pp_var.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/pp_var.json").read().strip().splitlines()[-1])
p=d["roofline_postprocess"]
print("$v", "frame", d["ms_per_step"], "pp_ms", p["kernel_ms"], "frac", p["frac"])
PY
done
python -m real_time_ray_tracer_b200.build --force > /dev/null 2>&1
