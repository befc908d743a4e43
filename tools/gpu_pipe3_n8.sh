#!/bin/bash
# 8-GPU box: the separate-GPU tests in both scheduling modes, then config 4 at N = 8 (three frames in flight),
# with and without the gather, and the full line (other configs, e2e) with the default gather
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "separate_gpus" 2>&1 | tail -3
RT_PIPELINE=1 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "separate_gpus" 2>&1 | tail -3
run() { # N tag args...
    N=$1; TAG=$2; shift 2
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_n${N}_$TAG.json 2> gpurun_out/bench_n${N}_$TAG.err || tail -20 gpurun_out/bench_n${N}_$TAG.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n${N}_$TAG.json").read().strip().splitlines()[-1])
    print("N=$N $TAG", d["config"]["workload"], "ms/step", d["ms_per_step"], "fps", d["frames_per_s"], "parity", {k:v for k,v in d.get("parity_check",{}).items() if k!="oracle_row_ranges"})
    for r in d["per_rank"]: print("  ", r)
    print("  e2e", (d.get("e2e") or {}).get("ms_per_step"))
    for k,v in d.get("other_configs",{}).items(): print("  ", k, v["ms_per_step"], v["frames_per_s"], [b[1] for b in v["bands"]])
except Exception as e: print("parse failed", e)
PY
}
LEAN="--no-cpu-baseline --no-e2e --no-other-configs"
run 8 none $LEAN --mode1-gather none
run 8 default --no-cpu-baseline
run 4 default $LEAN
