"""Hot instructions of one kernel from an .ncu-rep (source page): samples, dominant stall, SASS.

    python tools/ncu_hot.py file.ncu-rep kernel_name_substring [top_n] [launch_index]
"""
import csv
import io
import subprocess
import sys

rep, kname = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kname}",
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(io.StringIO(raw))]
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
data = []
for r in rows[hi + 1:]:
    if r and r[0] == "Kernel Name":      # the next captured launch: one is enough
        break
    if len(r) == len(hdr) and r[0].startswith("0x"):
        data.append(r)
print(rows[0][1][:100] if rows and len(rows[0]) > 1 else "")
ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[isamp]) for r in data)
print("instructions", len(data), "samples", tot, "warp-instructions executed", sum(int(r[iex]) for r in data))
agg = {s: sum(int(r[hdr.index(s)]) for r in data) for s in stalls}
print("stall totals:", [(k, v) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]])
top = sorted(enumerate(data), key=lambda ir: -int(ir[1][isamp]))[:top_n]
for i, r in sorted(top):
    best = max(stalls, key=lambda s: int(r[hdr.index(s)]))
    print(f"{i:5d} samp {int(r[isamp]):6d} ({100.0 * int(r[isamp]) / max(tot, 1):4.1f}%) exec {r[iex]:>8s} {best:22s} {r[ia].strip()[:80]}")
