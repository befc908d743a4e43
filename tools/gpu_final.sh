#!/bin/bash
# final one-GPU validation: smoke, the GPU suite in both scheduling modes, the bench line and the reference arm
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
bash tools/gpu_tests_both.sh
python bench.py > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/bench_final_ref.json 2> gpurun_out/bench_final_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_final_n1.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("metric","value","unit","ms_per_step","frames_per_s","gpu_launches","steps","warmup")})
print("roofline", d["roofline"]["frac"], d["roofline"]["kernel_ms"], "pp", d["roofline_postprocess"]["frac"], d["roofline_postprocess"]["kernel_ms"], d["roofline_postprocess"]["traffic"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "rgba8", d.get("e2e_rgba8",{}).get("ms_per_step"), "parity", d["parity_check"]["ok"], "clocks", d["clocks"])
print("cpu", d["cpu_baseline"])
for k,v in d["other_configs"].items(): print(k, v["ms_per_step"], v["roofline"]["frac"])
r=json.loads(open("gpurun_out/bench_final_ref.json").read().strip().splitlines()[-1])
print("ref", {k:r.get(k) for k in ("impl","value","unit","ms_per_step")}, r.get("cpu_baseline"), r["config"]==d["config"])
PY
