#!/bin/bash
# round 3: A/B of the plane-list / candidate-list unrolling switches of rt_trace_common.cuh (RT_PLANE_UNROLL,
# RT_MASKED_UNROLL): variants built beforehand as real_time_ray_tracer_b200/librt_v_<name>.so; each runs the headline
# workload and the other configs twice (order reversed the second time).  -> gpurun_out/r3_unroll_ab.txt
out=gpurun_out/r3_unroll_ab.txt
: > $out
names="base pl1 pl2 mk pl2mk"
for pass in 1 2; do
  [ $pass = 2 ] && names="pl2mk mk pl2 pl1 base"
  for n in $names; do
    RT_B200_LIB=$PWD/real_time_ray_tracer_b200/librt_v_$n.so python bench.py --no-e2e --no-cpu-baseline --no-parity-check --repeats 2 > gpurun_out/ab_$n.json 2>/dev/null
    python - $n $pass >> $out <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/ab_{sys.argv[1]}.json").read().strip().splitlines()[-1])
o = d.get("other_configs", {})
print(sys.argv[1], "pass", sys.argv[2], "config4 step", d["ms_per_step"], "trace", d["roofline"]["kernel_ms"],
      "spread", d["summary"].get("ms_per_step_min_median_max"),
      " | ".join(f"{k.split('_')[0]} {v['ms_per_step']} (k {v['roofline']['kernel_ms']})" for k, v in o.items()))
PY
  done
done
cat $out
