"""Per-view cost of the compositing kernels against the number of views in one launch (one wave vs several):
    python tools/views_scaling.py [n_gauss]
Same map, C copies of the SAME view, so the work per view is identical and only the schedule differs."""
import sys

import torch

sys.path.insert(0, ".")
from gslam_b200 import _lib, ops  # noqa: E402
from gslam_b200.rasterization import rasterization  # noqa: E402
from gslam_b200.synthetic import make_scene  # noqa: E402

n_gauss = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
g = make_scene(n_gauss, 1, "tum", 1.0).to("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for C in (1, 2, 3, 4, 8):
    vm, Ks = g.viewmats.expand(C, 4, 4).contiguous(), g.Ks.expand(C, 3, 3).contiguous()

    def run():
        ps = [q.clone().requires_grad_(True) for q in g.params()]
        out = rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, Ks, g.width, g.height, packed=False,
                            render_mode="RGB+D", log_uncertainties=ps[5])
        flush.zero_()
        out.render.sum().backward()
        flush.zero_()

    for _ in range(3):
        run()
    _lib.start_timing()
    for _ in range(5):
        run()
    t = _lib.stop_timing()
    print(C, {k.replace("gsb_", ""): round(v[1] * 1e3 / C, 1) for k, v in t.items()
              if k in ("gsb_composite_fwd", "gsb_composite_bwd", "gsb_segment_sort", "gsb_isect_scatter", "gsb_project_fwd", "gsb_project_bwd")},
          "us per view")
