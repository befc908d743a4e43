#!/bin/bash
# parity tests + one bench line per BASELINE.json config on one GPU -> gpurun_out/bench_<workload>.json
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log; tail -2 gpurun_out/pytest_gpu.log
run() { python bench.py --workload $1 --steps $2 --warmup $3 $4 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err || tail -5 gpurun_out/bench_$1.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$1.json").read().strip().splitlines()[-1])
    r=d["roofline"]; p=d.get("roofline_postprocess",{})
    print("$1", "ms/frame", d["ms_per_step"], "fps", d["frames_per_s"], "Mrays/s", d["value"], "frac", r["frac"], "trace_ms", r["kernel_ms"],
          "pp_ms", p.get("kernel_ms"), "pp_frac", p.get("frac"), "e2e", (d.get("e2e") or {}).get("ms_per_step"))
except Exception as e: print("$1 failed", e)
PY
}
run config2_ao_1080p 512 64
run config1_phong_640x480 512 64 --no-cpu-baseline
run config3_hybrid_1080p 120 16 --no-cpu-baseline
run config4_aopp_4k 64 16 --no-cpu-baseline
run config5_aopp_8k_aa16 16 4 "--no-cpu-baseline --no-e2e"
