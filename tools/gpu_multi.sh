#!/bin/bash
# tools/gpu_multi.sh N [bench args]: the separate-GPU tests (both scheduling modes) and the bench line on N GPUs
N=$1; shift
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "separate_gpus" 2>&1 | tail -3
RT_PIPELINE=1 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "separate_gpus" 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err || tail -20 gpurun_out/bench_n$N.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
    print("N=$N", d["config"]["workload"], "ms/step", d["ms_per_step"], "fps", d["frames_per_s"], "parity", d.get("parity_check"))
    for r in d["per_rank"]: print("  ", r)
    print("  e2e", (d.get("e2e") or {}).get("ms_per_step"), "clocks", d.get("clocks"))
    for k,v in d.get("other_configs",{}).items(): print("  ", k, v["ms_per_step"], v["frames_per_s"], v["bands"])
except Exception as e: print("parse failed", e)
PY
