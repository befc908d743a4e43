#!/bin/bash
# three frames in flight: the GPU suite in both scheduling modes, then the one-GPU bench line
bash tools/gpu_tests_both.sh
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_n1_pipe3.json 2> gpurun_out/bench_n1_pipe3.err || tail -5 gpurun_out/bench_n1_pipe3.err
python tools/band_pipeline_probe.py > gpurun_out/probe3.jsonl 2>/dev/null; cat gpurun_out/probe3.jsonl
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n1_pipe3.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline_postprocess"]["kernel_ms"], d["parity_check"]["ok"])
for k,v in d["other_configs"].items(): print(k, v["ms_per_step"])
PY
