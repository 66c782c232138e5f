#!/bin/bash
# A/B of the chunk-parallel backward settings: tools/ab_env.sh "CK SPLIT" "CK SPLIT" ...
for v in "$@"; do
  set -- $v; ck=$1; sp=$2
  GSB_CKPT_LEN=$ck GSB_BWD_SPLIT=$sp python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-mapping --no-sequence \
      > gpurun_out/abe_${ck}_${sp}.json 2> gpurun_out/abe_${ck}_${sp}.err || { echo "$v FAILED"; tail -3 gpurun_out/abe_${ck}_${sp}.err; continue; }
  python - "$ck" "$sp" <<'P'
import json, sys
ck, sp = sys.argv[1:3]
d = json.load(open(f"gpurun_out/abe_{ck}_{sp}.json"))
st = {s["kernel"].replace("gsb_", ""): round(s["ms"] * 1e3, 1) for s in d["roofline_stages"]}
print("ck", ck, "split", sp, "step_us", round(d["ms_per_step"] * 1e3, 1), st)
P
done
