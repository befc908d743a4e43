#!/bin/bash
# rectangle / big-scene GPU tests, the golden fixtures, then a lean bench line to compare with the previous build
python -m pytest tests/test_rectangles_big_scenes.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -15
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_n1_rect.json 2> gpurun_out/bench_n1_rect.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n1_rect.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline_postprocess"]["kernel_ms"], d["parity_check"]["ok"])
for k,v in d["other_configs"].items(): print(k, v["ms_per_step"])
PY
