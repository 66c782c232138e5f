#!/bin/bash
# A/B of kernel variants on the GPU box: tools/ab.sh name1 name2 ...  ("base" = the product library)
# -> gpurun_out/ab_<name>.json, and one summary line per variant (stage times in us)
for v in "$@"; do
  if [ "$v" = base ]; then lib=gslam_b200/libgsb.so; else lib=build/var_$v/libgsb_$v.so; fi
  GSB_LIB=$lib python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-mapping --no-sequence \
      > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err || { echo "$v FAILED"; tail -3 gpurun_out/ab_$v.err; continue; }
  python - "$v" <<'P'
import json, sys
v = sys.argv[1]
d = json.load(open(f"gpurun_out/ab_{v}.json"))
st = {s["kernel"].replace("gsb_", ""): round(s["ms"] * 1e3, 1) for s in d["roofline_stages"]}
print(v, "step_us", round(d["ms_per_step"] * 1e3, 1), st)
P
done
