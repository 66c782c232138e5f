"""MappingStep on the CUDA kernels vs the same step with the oracle injected (CPU)."""
import pytest
import torch

from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep
from gslam_b200.optim import FusedAdam
from gslam_b200.synthetic import make_scene, target_images
from oracle import torch_ref as R

pytestmark = pytest.mark.gpu


def _oracle_render(*a, **kw):
    out = R.rasterization(*a, **kw)
    out.render = torch.cat([out.rgbs, out.depthmaps[..., None], out.betas[..., None]], dim=-1)
    return out


def _oracle_loss(render, gt, exposure, w, mode):
    photo = R.mapping_photometric_loss(R.exposure_affine(render[..., :3], exposure), gt, render[..., 4])
    ssim = 1.0 - R.fused_ssim(render[..., :3].permute(0, 3, 1, 2), gt.permute(0, 3, 1, 2), padding="valid")
    return (1 - w) * photo + w * ssim, photo.detach()


class _RefAdam:
    def __init__(self, params):
        self.opts = [torch.optim.Adam([p], lr=DEFAULT_LRS[n]) for p, n in zip(params, PARAM_NAMES)]

    def step(self, decay_param=None, vis_count=None, decay=0.995):
        for o in self.opts:
            o.step()
        with torch.no_grad():
            decay_param[vis_count > 1] *= decay


def test_mapping_step_matches_oracle_step(cuda):
    sc = make_scene(1500, 3, "tum", 0.2)
    gt = target_images(3, sc.height, sc.width)
    exposure = torch.tensor([[0.02, 0.01], [-0.01, 0.0], [0.0, -0.02]])
    ref = FlatSplats(sc.params())
    ref_step = MappingStep(ref, _RefAdam(ref.params), render_fn=_oracle_render, loss_fn=_oracle_loss,
                           depth_tv_weight=1e-6)
    g = sc.to(cuda)
    ours = FlatSplats(g.params())
    opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[n]) for p, n in zip(ours.params, PARAM_NAMES)])
    our_step = MappingStep(ours, opt, depth_tv_weight=1e-6)
    for it in range(2):
        r0 = ref_step(sc.viewmats, sc.Ks, gt, exposure, sc.width, sc.height, 3)
        r1 = our_step(g.viewmats, g.Ks, gt.to(cuda), exposure.to(cuda), sc.width, sc.height, 3)
        assert abs(float(r1.total) - float(r0.total)) < 1e-4 * abs(float(r0.total))
        assert abs(float(r1.photometric) - float(r0.photometric)) < 1e-4 * abs(float(r0.photometric))
        gr, go = ref.flat_grad(), ours.flat_grad().cpu()
        assert float((go - gr).norm() / gr.norm()) < 1e-3
        assert (r1.vis_count.cpu() != r0.vis_count).float().mean() < 1e-3
    # Adam's first steps move every coordinate by ~lr regardless of the gradient's size, so compare
    # the parameter displacement, not the parameters
    p0 = FlatSplats(sc.params()).flat
    dr, do = ref.flat.detach() - p0, ours.flat.detach().cpu() - p0
    assert float((do - dr).norm() / dr.norm()) < 2e-2


def test_visibility_count_and_isotropic_reg_kernels_match_torch(cuda):
    """gsb_vis_count / gsb_isotropic_reg against the torch expressions of gslam/backend.py:287-296."""
    from gslam_b200.mapping import add_isotropic_grad_, visibility_count
    g = torch.Generator().manual_seed(0)
    n = 10_007
    radii = (torch.randint(-2, 6, (5, n), generator=g) * (torch.rand(5, n, generator=g) > 0.4)).to(torch.int32)
    vc = visibility_count(radii.to(cuda))
    assert vc.dtype == torch.int32 and torch.equal(vc.cpu(), (radii > 0).sum(0).to(torch.int32))
    s = (-4.0 + 0.7 * torch.randn(n, 3, generator=g))
    grad0 = torch.randn(n, 3, generator=g)
    ref_grad = grad0.clone()
    ref_loss = add_isotropic_grad_(ref_grad, s, (radii > 0).sum(0) > 0, 5e-4)         # CPU: torch expression
    gpu_grad = grad0.clone().to(cuda)
    gpu_loss = add_isotropic_grad_(gpu_grad, s.to(cuda), vc, 5e-4)                     # CUDA: one kernel
    assert torch.allclose(gpu_grad.cpu(), ref_grad, rtol=1e-5, atol=1e-7)
    assert abs(float(gpu_loss) - float(ref_loss)) < 1e-4 * abs(float(ref_loss))
