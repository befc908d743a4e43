#!/bin/bash
# tools/gpu_multi_bench.sh "N N ..." [bench args]: the bench line on N GPUs of this box, for each N
NS=$1; shift
for N in $NS; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err || tail -20 gpurun_out/bench_n$N.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
    print("N=$N", d["config"]["workload"], "ms/step", d["ms_per_step"], "fps", d["frames_per_s"], "parity", {k:v for k,v in d.get("parity_check",{}).items() if k!="oracle_row_ranges"})
    for r in d["per_rank"]: print("  ", r)
    print("  e2e", (d.get("e2e") or {}).get("ms_per_step"), "clocks", d.get("clocks"))
    for k,v in d.get("other_configs",{}).items(): print("  ", k, v["ms_per_step"], v["frames_per_s"], [b[1] for b in v["bands"]])
except Exception as e: print("parse failed", e)
PY
done
