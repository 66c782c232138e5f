"""Generate tests/golden/raster_small.npz from the CPU oracle (oracle/torch_ref.py).

The reference's kernels live in un-vendored CUDA extensions (gsplat fork, fused-ssim)
that cannot be imported in this container, and it ships no fixtures of its own
(SURVEY.md 8c), so these vectors pin the ORACLE's behaviour: a later change to the
oracle or to the kernels that moves any of them is caught by tests/test_golden.py.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gslam_b200.synthetic import make_scene, target_images  # noqa: E402
from oracle import torch_ref as R  # noqa: E402

N, C, SCALE = 400, 2, 0.1      # 64 x 48 pixels, 4 x 3 tiles per view


def build():
    torch.manual_seed(0)
    torch.set_num_threads(1)
    sc = make_scene(N, C, "tum", SCALE, log_scale_mean=float(np.log(0.03)))
    gt = target_images(C, sc.height, sc.width)
    exposure = torch.tensor([[0.05, -0.02], [-0.03, 0.01]])
    ps = [p.clone().requires_grad_(True) for p in sc.params()]
    vm = sc.viewmats.clone().requires_grad_(True)
    ex = exposure.clone().requires_grad_(True)
    out = R.rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, sc.Ks, sc.width, sc.height,
                          backgrounds=torch.tensor([[0.1, 0.2, 0.3]] * C), render_mode="RGB+D",
                          log_uncertainties=ps[5])
    total, photo = R.mapping_loss(out, gt, ex, ps[2])
    total.backward()
    d = dict(
        rgbs=out.rgbs, alphas=out.alphas, depthmaps=out.depthmaps, betas=out.betas, radii=out.radii,
        means2d=out.means2d, depths=out.depths, conics=out.conics, tiles_per_gauss=out.tiles_per_gauss,
        isect_ids=out.isect_ids, flatten_ids=out.flatten_ids, isect_offsets=out.isect_offsets,
        n_touched=out.n_touched, loss_total=total, loss_photo=photo,
        g_means=ps[0].grad, g_quats=ps[1].grad, g_scales=ps[2].grad, g_opacities=ps[3].grad,
        g_colors=ps[4].grad, g_log_unc=ps[5].grad, g_viewmats=vm.grad, g_exposure=ex.grad,
        gt=gt, exposure=exposure)
    return {k: v.detach().numpy() for k, v in d.items()}


if __name__ == "__main__":
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "raster_small.npz")
    np.savez_compressed(path, **build())
    print("wrote", path, os.path.getsize(path), "bytes")
