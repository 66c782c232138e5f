import json,sys
src=open(sys.argv[1]) if len(sys.argv)>1 else sys.stdin
d=json.loads(src.read().strip().splitlines()[-1])
print("value", round(d["value"],1), "ms", round(d["ms_per_step"],4), "eager_ms", round(d["eager_ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "launches/step", d["launch_calls_per_step"])
tot=0
for s in d["roofline_stages"]:
    tot+=s["ms"]; print(f'{s["kernel"]:22s} {s["ms"]*1000:8.1f} us  hbm_frac {s["frac"]:.3f}', s.get("fp32",""))
print("sum kernels us", round(tot*1000,1))

if d.get("mapping"): print("mapping", round(d["mapping"]["value"],1), "views/s", round(d["mapping"]["ms_per_step"],4), "ms")
if d.get("tracking_closure"): print("tracking_closure", round(d["tracking_closure"]["value"],1), "it/s", round(d["tracking_closure"]["ms_per_step"],4), "ms", d["tracking_closure"]["launch_calls_per_step"], "launches")
