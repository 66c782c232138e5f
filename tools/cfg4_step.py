"""BASELINE.json configs[3] shape on ONE rank: 3 M Gaussians, SH degree 3, 1280x720 (OAK-D x 1/3),
2 views per GPU (16 views over 8 GPUs) -- render RGB+D+beta, photometric + SSIM loss, backward to
every Gaussian parameter (SH coefficients included).  Prints per-entry-point CUDA-event times.

    python tools/cfg4_step.py [n_gauss] [views_per_rank]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gslam_b200 import _lib  # noqa: E402
from gslam_b200.losses import photometric_ssim_loss  # noqa: E402
from gslam_b200.rasterization import rasterization  # noqa: E402
from gslam_b200.synthetic import make_scene, target_images  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
views = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda:0")
sc = make_scene(n, views, "oakd", 1.0).to(dev)
g = torch.Generator().manual_seed(4)
sh = (0.1 * torch.randn(n, 16, 3, generator=g))
sh[:, 0] = (torch.sigmoid(sc.logit_colors.cpu()) - 0.5) / 0.28209479177387814
sh = sh.to(dev).requires_grad_(True)
ps = [p.clone().requires_grad_(True) for p in (sc.means, sc.quats, sc.log_scales, sc.logit_opacities)]
unc = sc.log_uncertainties.clone().requires_grad_(True)
gt = target_images(views, sc.height, sc.width).to(dev)
exposure = torch.zeros(views, 2, device=dev)
bg = torch.zeros(views, 3, device=dev)


def step():
    for p in ps + [sh, unc]:
        p.grad = None
    out = rasterization(ps[0], ps[1], ps[2], ps[3], sh, sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                        render_mode="RGB+D", log_uncertainties=unc, backgrounds=bg, sh_degree=3)
    total, _ = photometric_ssim_loss(out.render, gt, exposure, ssim_weight=0.2)
    total.backward()
    return out, total


for _ in range(2):
    out, total = step()
torch.cuda.synchronize()
print(f"N={n} views={views} {sc.width}x{sc.height} SH3: isects {out.isect_ids.numel()}, loss {float(total):.5f}, "
      f"visible {int((out.radii > 0).sum())}, peak mem {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(5):
    step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"step {ms:.3f} ms device ({(time.perf_counter() - t0) * 200:.3f} ms wall) -> {views * 1e3 / ms:.1f} views/s per GPU")
_lib.start_timing()
for _ in range(3):
    step()
kt = _lib.stop_timing()
tot = 0.0
for name, (calls, msk) in sorted(kt.items(), key=lambda kv: -kv[1][1] * kv[1][0]):
    print(f"{name:24s} calls/step {calls / 3:4.1f}  {msk * 1e3:9.1f} us/call")
    tot += msk * calls / 3
print(f"sum of C-ABI calls: {tot:.3f} ms/step")
assert all(torch.isfinite(p.grad).all() for p in ps + [sh, unc])
