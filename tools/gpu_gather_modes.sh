#!/bin/bash
# tools/gpu_gather_modes.sh: config 4 on 8 GPUs with the three ways of bringing the bands to rank 0 (float4 copies,
# RGBA8 copies, none), then the RGBA8 line at N = 4 and 2.  One JSON line per run under gpurun_out/.
run() { # N tag args...
    N=$1; TAG=$2; shift 2
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_n${N}_$TAG.json 2> gpurun_out/bench_n${N}_$TAG.err || tail -20 gpurun_out/bench_n${N}_$TAG.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n${N}_$TAG.json").read().strip().splitlines()[-1])
    print("N=$N $TAG", d["config"]["workload"], "ms/step", d["ms_per_step"], "fps", d["frames_per_s"], "parity", {k:v for k,v in d.get("parity_check",{}).items() if k!="oracle_row_ranges"})
    for r in d["per_rank"]: print("  ", r)
    for k,v in d.get("other_configs",{}).items(): print("  ", k, v["ms_per_step"], v["frames_per_s"], [b[1] for b in v["bands"]])
except Exception as e: print("parse failed", e)
PY
}
LEAN="--no-cpu-baseline --no-e2e --no-other-configs"
run 8 none $LEAN --mode1-gather none
run 8 copy $LEAN --mode1-gather copy
run 8 copy8 --no-cpu-baseline --no-e2e --mode1-gather copy8
run 4 copy8 $LEAN --mode1-gather copy8
run 2 copy8 $LEAN --mode1-gather copy8
