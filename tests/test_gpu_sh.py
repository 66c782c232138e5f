"""GPU parity of the spherical-harmonics colour extension (SURVEY 8a "SH note") against the oracle.

Not a reference feature: with sh_degree=None the behaviour is the reference's sigmoid path (covered
by test_gpu_parity.py); here the oracle's sh_colors is the checker, images 1e-4 / gradients 1e-3.
"""
import pytest
import torch

from gslam_b200.rasterization import rasterization
from gslam_b200.synthetic import make_scene
from oracle import torch_ref as R

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-12))


def _coeffs(n, K, seed=11):
    g = torch.Generator().manual_seed(seed)
    sh = 0.3 * torch.randn(n, K, 3, generator=g)
    sh[:, 0] += 0.5 * torch.randn(n, 3, generator=g)      # DC term: some colours clamp at 0
    return sh


@pytest.mark.parametrize("degree,K", [(0, 1), (1, 4), (2, 9), (3, 16), (2, 16)])
@pytest.mark.parametrize("n,c", [(2500, 2)])
def test_sh_render_and_gradients(cuda, degree, K, n, c):
    sc = make_scene(n, c, "tum", 0.25)
    sh = _coeffs(n, K)

    def run(dev, fn):
        ps = [p.clone().to(dev).requires_grad_(True) for p in
              (sc.means, sc.quats, sc.log_scales, sc.logit_opacities, sh, sc.log_uncertainties)]
        vm = sc.viewmats.clone().to(dev).requires_grad_(True)
        out = fn(ps[0], ps[1], ps[2], ps[3], ps[4], vm, sc.Ks.to(dev), sc.width, sc.height, packed=False,
                 render_mode="RGB+D", log_uncertainties=ps[5], backgrounds=torch.zeros(c, 3, device=dev),
                 sh_degree=degree)
        return out, ps, vm

    ref, rps, _ = run("cpu", R.rasterization)
    out, gps, _ = run(cuda, rasterization)
    assert rel_err(out.rgbs, ref.rgbs) < 1e-4
    assert rel_err(out.alphas, ref.alphas) < 1e-4
    w = torch.rand(ref.rgbs.shape, generator=torch.Generator().manual_seed(5))
    (ref.rgbs * w).sum().backward()
    (out.rgbs * w.to(cuda)).sum().backward()
    names = ["means", "quats", "log_scales", "logit_opacities", "sh", "log_uncertainties"]
    for name, p, gp in zip(names, rps, gps):
        if name == "log_uncertainties":
            continue    # the loss does not touch beta
        assert rel_err(gp.grad, p.grad) < 1e-3, (name, rel_err(gp.grad, p.grad))
    # coefficients above the active degree receive exactly zero
    Ka = (degree + 1) ** 2
    assert bool((gps[4].grad[:, Ka:] == 0).all())


def test_sh_degree0_is_view_independent_and_none_is_sigmoid(cuda):
    sc = make_scene(2000, 2, "tum", 0.25).to(cuda)
    p = sc.params()
    bg = torch.zeros(2, 3, device=cuda)
    base = rasterization(p[0], p[1], p[2], p[3], p[4], sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                         backgrounds=bg)
    # degree 0 with c0 = (sigmoid(logit) - 0.5) / Y0 reproduces the sigmoid colours
    c0 = ((torch.sigmoid(p[4]) - 0.5) / 0.28209479177387814)[:, None, :]
    out = rasterization(p[0], p[1], p[2], p[3], c0, sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                        backgrounds=bg, sh_degree=0)
    assert rel_err(out.rgbs, base.rgbs) < 1e-5
    with pytest.raises(ValueError):
        rasterization(p[0], p[1], p[2], p[3], c0, sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                      backgrounds=bg, sh_degree=2)
    with pytest.raises(NotImplementedError):
        rasterization(p[0], p[1], p[2], p[3], c0, sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                      backgrounds=bg)      # [N,1,3] colours without sh_degree: not the reference surface
