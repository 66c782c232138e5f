"""Instruction counts per kernel from the built library (no GPU needed):
    python tools/sass_counts.py [gslam_b200/libgsb.so] > profiles/r2_sass_counts.txt
What the kernels are made of: cp.async gathers (LDGSTS), SFU ops, shared / global atomics, NVLink-multicast
accesses (multimem.* -> LDGMC / ST..MC / RED..MC), 16-byte shared loads."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "gslam_b200/libgsb.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
classes = [("LDGSTS (cp.async)", r"^LDGSTS"), ("MUFU.EX2", r"^MUFU\.EX2"), ("MUFU.RCP", r"^MUFU\.RCP"),
           ("shared atomics (ATOMS)", r"^ATOMS"), ("global atomics (ATOMG / ATOM / RED)", r"^(ATOMG|ATOM\.|RED)"),
           ("multicast (multimem: *MC*)", r"MC\b|^LDGMC|\.MC\."), ("LDS.128", r"^LDS(\.U)?\.128"),
           ("FFMA / FMUL / FADD", r"^(FFMA|FMUL|FADD)"), ("warp votes / shuffles", r"^(VOTE|SHFL|REDUX|MATCH)")]
counts = {c: collections.Counter() for c, _ in classes}
total = collections.Counter()
fn = None
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if not m or fn is None:
        continue
    op = m.group(1)
    total[fn] += 1
    for c, pat in classes:
        if re.search(pat, op):
            counts[c][fn] += 1
print(f"# SASS instruction counts per kernel of {lib} (sm_100a), static counts, `python tools/sass_counts.py`")
for c, _ in classes:
    print(f"\n## {c}")
    for f, n in counts[c].most_common():
        print(f"{n:7d}  {f}")
print("\n## total instructions")
for f, n in total.most_common():
    print(f"{n:7d}  {f}")
