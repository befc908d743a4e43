#!/bin/bash
# multi-GPU bench lines: tools/gpu_scale.sh N workload steps warmup [extra bench flags]
N=$1; WL=$2; K=${3:-256}; W=${4:-32}; EXTRA=$5; TAG=${6:-direct}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $WL --steps $K --warmup $W --no-cpu-baseline $EXTRA > gpurun_out/scale_${WL}_n${N}_$TAG.json 2> gpurun_out/scale_${WL}_n${N}_$TAG.err || tail -20 gpurun_out/scale_${WL}_n${N}_$TAG.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_${WL}_n${N}_$TAG.json").read().strip().splitlines()[-1])
    print("$WL n=$N $TAG ms/frame", d["ms_per_step"], "fps", d["frames_per_s"], "Mrays/s", d["value"], "bands", d["config"]["bands"], "e2e", (d.get("e2e") or {}).get("ms_per_step"))
except Exception as e: print("$WL n=$N $TAG failed", e)
PY
