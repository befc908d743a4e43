#!/bin/bash
# both arms under torchrun at N = 2, as the driver launches them
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $T bench.py --impl reference --gpus 2 --steps 4 --warmup 1 > gpurun_out/n2_ref.json 2> gpurun_out/n2_ref.err; echo "ref rc=$?"
timeout 900 $T bench.py --gpus 2 > gpurun_out/n2_b200.json 2> gpurun_out/n2_b200.err; echo "b200 rc=$?"
python - <<PY
import json
r=json.loads(open("gpurun_out/n2_ref.json").read().strip().splitlines()[-1])
d=json.loads(open("gpurun_out/n2_b200.json").read().strip().splitlines()[-1])
print("ref", {k:r.get(k) for k in ("impl","value","unit","n_gpus","ms_per_step")}, r.get("cpu_baseline"), "same config", r["config"]==d["config"])
print("b200", {k:d.get(k) for k in ("value","ms_per_step","frames_per_s","n_gpus","scaling","gpu_launches")}, "parity", {k:v for k,v in d["parity_check"].items() if k!="oracle_row_ranges"})
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "clocks", d["clocks"])
for x in d["per_rank"]: print("  ", x)
for k,v in d["other_configs"].items(): print("  ", k, v["ms_per_step"])
PY
