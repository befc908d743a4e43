#!/bin/bash
# A/B of prebuilt library variants (real_time_ray_tracer_b200/librt_b200_<tag>.so, built with RT_NVCC_EXTRA=-D...):
#   tools/gpu_variants2.sh "<tag> ..." "<workload:steps:warmup> ..."      ("default" = the product library)
for tag in $1; do
  lib=""; [ $tag != default ] && lib=$PWD/real_time_ray_tracer_b200/librt_b200_$tag.so
  for spec in $2; do
    IFS=: read wl st wu <<< "$spec"
    RT_B200_LIB=$lib python bench.py --workload $wl --steps $st --warmup $wu --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check > gpurun_out/var.json 2> gpurun_out/var.err || tail -3 gpurun_out/var.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/var.json").read().strip().splitlines()[-1])
    p=d.get("roofline_postprocess",{})
    print("$tag $wl", "frame", d["ms_per_step"], "trace_ms", d["roofline"]["kernel_ms"], "frac", d["roofline"]["frac"], "pp_ms", p.get("kernel_ms"), "pp_frac", p.get("frac"), "Mrays/s", d["value"], flush=True)
except Exception as e: print("$tag $wl failed", e)
PY
  done
done
