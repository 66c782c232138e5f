"""The binning kernels in isolation (tile scan, scatter, segment sort) on a projected scene:
    python tools/bin_bench.py [n_views] [n_gauss]      (GSB_LIB=... for a variant library)
CUDA-event times per call, L2 flushed before every call."""
import math
import sys

import torch

sys.path.insert(0, ".")
from gslam_b200 import ops  # noqa: E402
from gslam_b200._lib import call, ptr, stream  # noqa: E402
from gslam_b200.synthetic import make_scene  # noqa: E402

C = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 500_000
g = make_scene(n, C, "tum", 1.0).to("cuda")
tw, th = math.ceil(g.width / 16), math.ceil(g.height / 16)
with torch.no_grad():
    radii, m2, rec, tiles, counts = ops.project(g.means, g.quats, g.log_scales, g.logit_opacities, g.logit_colors,
                                                g.log_uncertainties, g.viewmats, g.Ks, g.width, g.height)
T = tw * th
dev = radii.device
offsets = torch.empty(C * T + 1, dtype=torch.int32, device=dev)
cursors = torch.empty(C * T * ops.CSTRIDE, dtype=torch.int32, device=dev)
n_dev = torch.empty(1, dtype=torch.int64, device=dev)
overflow = torch.zeros(1, dtype=torch.int32, device=dev)
call("gsb_tile_scan", ptr(counts), C, tw, th, ptr(offsets), ptr(cursors), ptr(n_dev), stream())
cap = int(n_dev.item())
depth_id = torch.empty(cap, dtype=torch.int64, device=dev)
isect_ids = torch.empty(cap, dtype=torch.int64, device=dev)
flatten_ids = torch.empty(cap, dtype=torch.int32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
tile_n_bits, _ = ops.tile_bits(tw, th, C)
times = {"scan": [], "scatter": [], "sort": []}


def timed(name, fn):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record()
    torch.cuda.synchronize()
    times[name].append(e0.elapsed_time(e1) * 1e3)


for it in range(6):
    timed("scan", lambda: call("gsb_tile_scan", ptr(counts), C, tw, th, ptr(offsets), ptr(cursors), ptr(n_dev), stream()))
    timed("scatter", lambda: call("gsb_isect_scatter", ptr(m2), ptr(radii), ptr(rec), C, n, tw, th, ptr(cursors), cap,
                                  ptr(depth_id), ptr(overflow), stream()))
    if "nosort" not in sys.argv:
        timed("sort", lambda: call("gsb_segment_sort", ptr(depth_id), ptr(isect_ids), ptr(offsets), C, tw, th, tile_n_bits,
                                   cap, ptr(isect_ids), ptr(flatten_ids), stream()))
print(f"C={C} N={n} I={cap} ({cap / (C * T):.0f} per tile):",
      {k: round(sorted(v)[len(v) // 2], 1) for k, v in times.items() if v}, "us (median of 6)")
