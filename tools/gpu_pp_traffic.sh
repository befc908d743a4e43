#!/bin/bash
# Post-processing kernel at config 4, per library variant (real_time_ray_tracer_b200/librt_b200_<tag>.so, built with
# RT_NVCC_EXTRA=-D...): the bench's kernel time and HBM fraction, then DRAM bytes of three launches under ncu.
#   tools/gpu_pp_traffic.sh "<tag> ..."      ("default" = the product library)
for tag in $1; do
  lib=""; [ $tag != default ] && lib=$PWD/real_time_ray_tracer_b200/librt_b200_$tag.so
  CMD="python bench.py --workload config4_aopp_4k --steps 64 --warmup 16 --no-e2e --no-cpu-baseline --no-other-configs --no-parity-check"
  RT_B200_LIB=$lib $CMD > gpurun_out/var.json 2> gpurun_out/var.err || { tail -3 gpurun_out/var.err; continue; }
  RT_B200_LIB=$lib ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:rt_ao_postprocess -s 40 -c 3 --csv --log-file gpurun_out/pp_traffic_$tag.csv $CMD > /dev/null 2>&1
  python - <<PY
import json, csv
d=json.loads(open("gpurun_out/var.json").read().strip().splitlines()[-1])
p=d["roofline_postprocess"]
rows=list(csv.DictReader(l for l in open("gpurun_out/pp_traffic_$tag.csv") if l.startswith('"')))
m={}
for r in rows: m.setdefault(r["Metric Name"],[]).append((float(r["Metric Value"].replace(",","")), r["Metric Unit"]))
def avg(k):
    v=m.get(k,[]); 
    if not v: return None
    mul={"byte":1,"Kbyte":1e3,"Mbyte":1e6,"Gbyte":1e9,"ns":1,"us":1e3,"usecond":1e3,"nsecond":1}.get(v[0][1],1)
    return sum(x for x,_ in v)/len(v)*mul
rd,wr,t=avg("dram__bytes_read.sum"),avg("dram__bytes_write.sum"),avg("gpu__time_duration.sum")
print("%-12s frame %.4f ms  pp %.4f ms  frac %.4f  alg %.1f MB | ncu: read %.1f MB write %.1f MB  %.1f us  traffic/alg %.3f" % ("$tag", d["ms_per_step"], p["kernel_ms"], p["frac"], p["algorithmic_bytes"]/1e6, rd/1e6, wr/1e6, t/1e3, (rd+wr)/p["algorithmic_bytes"]), flush=True)
PY
done
