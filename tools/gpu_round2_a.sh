#!/bin/bash
# round 2, first GPU call: parity tests, the default bench line (both arms), launch list, band sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err || tail -20 gpurun_out/bench_default.err
timeout 300 python bench.py --impl reference --steps 8 --warmup 3 --ref-budget 30 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err || tail -5 gpurun_out/bench_reference.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_config4.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check > gpurun_out/ncu_launch.log 2>&1
timeout 600 python tools/band_sweep.py > gpurun_out/band_sweep.jsonl 2> gpurun_out/band_sweep.err || tail -5 gpurun_out/band_sweep.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","ms_per_step","frames_per_s","parity_check","clocks")})
    print("roofline", d["roofline"]["frac"], d["roofline"]["kernel_ms"], "pp", d["roofline_postprocess"]["frac"], d["roofline_postprocess"]["kernel_ms"])
    print("e2e", d.get("e2e",{}).get("ms_per_step"), "cpu", d.get("cpu_baseline"))
    for k,v in d.get("other_configs",{}).items(): print(k, v["ms_per_step"], v["value"], v["roofline"]["frac"], v.get("roofline_postprocess",{}).get("frac"))
except Exception as e: print("bench parse failed", e)
PY
cat gpurun_out/band_sweep.jsonl
