#!/bin/bash
# 8 GPUs: the single-pass modes (configs 1-3 as other_configs legs) with the band copies instead of fused stores,
# then config 2 alone with copies and with no gather at all (compute only)
run() { # N tag args...
    N=$1; TAG=$2; shift 2
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_n${N}_$TAG.json 2> gpurun_out/bench_n${N}_$TAG.err || tail -20 gpurun_out/bench_n${N}_$TAG.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n${N}_$TAG.json").read().strip().splitlines()[-1])
    print("N=$N $TAG", d["config"]["workload"], "ms/step", d["ms_per_step"], "fps", d["frames_per_s"], "parity", {k:v for k,v in d.get("parity_check",{}).items() if k!="oracle_row_ranges"})
    for r in d["per_rank"]: print("  ", r)
    for k,v in d.get("other_configs",{}).items(): print("  ", k, v["ms_per_step"], v["frames_per_s"], v["roofline"]["kernel_ms"])
except Exception as e: print("parse failed", e)
PY
}
run 8 spcopy --no-cpu-baseline --no-e2e --single-pass-gather copy
run 8 c2none --no-cpu-baseline --no-e2e --no-other-configs --workload config2_ao_1080p --single-pass-gather none
run 8 c2store --no-cpu-baseline --no-e2e --no-other-configs --workload config2_ao_1080p --single-pass-gather store
