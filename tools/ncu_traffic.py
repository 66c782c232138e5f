"""profiles/traffic.json from an `ncu --set full` capture: DRAM bytes (read + write) per launch of every
kernel, keyed by the C-ABI entry point that launches it (bench.py puts them in `roofline.traffic`).

    python tools/ncu_traffic.py gpurun_out/prof.ncu-rep profiles/traffic.json
"""
import csv
import datetime
import io
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import csrc_sha  # noqa: E402  (the hash bench.py checks before quoting these numbers)

ENTRY = {"project_fwd_kernel": "gsb_project_fwd", "project_bwd_kernel": "gsb_project_bwd",
         "tile_scan_kernel": "gsb_tile_scan", "isect_scatter": "gsb_isect_scatter",
         "segment_sort_kernel": "gsb_segment_sort", "composite_fwd_kernel": "gsb_composite_fwd",
         "composite_bwd_kernel": "gsb_composite_bwd", "photo_loss_kernel": "gsb_photo_loss",
         "scale_kernel": "gsb_scale_by_scalar", "ssim_fwd_kernel": "gsb_ssim_fwd", "ssim_bwd_kernel": "gsb_ssim_bwd",
         "adam_multi_kernel": "gsb_adam_multi"}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
kn, rd, wr, du = (hdr.index(c) for c in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
acc = {}
for r in rows[2:]:
    key = next((v for k, v in ENTRY.items() if k in r[kn]), None)
    if key is None:
        continue
    b = float(r[rd]) * UNIT[units[rd]] + float(r[wr]) * UNIT[units[wr]]
    acc.setdefault(key, []).append((b, float(r[du])))
res = {"source": rep.split("/")[-1], "csrc_sha": csrc_sha(),
       "captured": datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%dT%H:%MZ"), "note": "mean over the captured launches; ncu serialises kernels and flushes caches, so "
       "inputs that the previous kernel left in L2 during a real step are re-read from DRAM here",
       "dram_bytes_per_launch": {k: round(sum(x[0] for x in v) / len(v)) for k, v in acc.items()},
       "ncu_us_per_launch": {k: round(sum(x[1] for x in v) / len(v), 2) for k, v in acc.items()},
       "launches_captured": {k: len(v) for k, v in acc.items()}}
with open(out, "w") as f:
    json.dump(res, f, indent=1)
print(json.dumps(res, indent=1))
