#!/bin/bash
# the default bench line at N = 8 as the driver launches it
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 > gpurun_out/n8_final.json 2> gpurun_out/n8_final.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/n8_final.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_step","frames_per_s","n_gpus","gpu_launches")}, "parity", {k:v for k,v in d["parity_check"].items() if k!="oracle_row_ranges"})
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "clocks", d["clocks"], "gather", d["config"].get("gather"))
for x in d["per_rank"]: print("  ", x)
for k,v in d["other_configs"].items(): print("  ", k, v["ms_per_step"], v["frames_per_s"])
PY
