#!/bin/bash
# A/B of the AO kernels on the GPU box: parity tests, then short benches per kernel
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for k in ${KERNELS:-phased tile}; do
  RT_AO_KERNEL=$k python bench.py --steps 256 --warmup 32 --no-cpu-baseline --no-e2e > gpurun_out/bench_c2_$k.log 2>&1
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_c2_$k.log").read().strip().splitlines()[-1])
    print("$k c2", d["ms_per_step"], d["value"], d["roofline"]["frac"], d["roofline"]["kernel_ms"])
except Exception as e:
    print("$k c2 failed", e); print(open("gpurun_out/bench_c2_$k.log").read()[-2000:])
PY
done
