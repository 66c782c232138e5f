#!/bin/bash
# The GPU evidence of a round in one gpurun call: full GPU suite, launch list, one --set full capture summarised ON THE BOX
# (the .ncu-rep stays there: gpurun_out/ is capped at 64 MiB), traffic.json, the bench line and the reference arm.
mkdir -p gpurun_out
B="--steps 2 --warmup 3 --no-cpu-baseline --no-mapping --no-sequence --no-cfg1"
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_gpu_tests_final.log; tail -2 gpurun_out/r2_gpu_tests_final.log
python bench.py $B > gpurun_out/pre.json 2> gpurun_out/pre.err || { echo "bench failed"; tail -5 gpurun_out/pre.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py $B > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/r2_launches.csv > gpurun_out/r2_launches_summary.csv
ncu --set full --clock-control none --import-source on -c 60 -o /tmp/prof_r2 python bench.py $B > gpurun_out/ncu_f.log 2>&1
python tools/ncu_summary.py /tmp/prof_r2.ncu-rep > gpurun_out/r2_ncu_full_summary.txt
python tools/ncu_hot.py /tmp/prof_r2.ncu-rep composite_fwd 25 > gpurun_out/r2_hot_composite_fwd.txt
python tools/ncu_hot.py /tmp/prof_r2.ncu-rep composite_bwd 25 > gpurun_out/r2_hot_composite_bwd.txt
python tools/ncu_traffic.py /tmp/prof_r2.ncu-rep gpurun_out/traffic.json && cp gpurun_out/traffic.json profiles/traffic.json
python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/ref.err; echo "ref rc=$?"
python - <<'P'
import json
d = json.loads(open("gpurun_out/r2_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step")}, "e2e", d["e2e"]["value"], d["e2e"].get("h2d_copy_ms"), "serial", d["e2e"]["serial_value"])
print("roofline", d["roofline"]["kernel"], d["roofline"]["frac"], "traffic", d["roofline"]["traffic"], d.get("traffic_source"))
print("parity ok", d["parity"]["ok"], "mapping", d["mapping"]["value"], "seq", (d.get("sequence") or {}).get("rgbd", {}).get("value"))
P
