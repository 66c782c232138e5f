"""MappingStep on the CUDA kernels (gradient rows written by gsb_project_bwd, gsb_map_update) vs the same
step with the oracle renderer / loss and the torch restatement of the update injected (tests/cpu_stubs.py)."""
import pytest
import torch

from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep
from gslam_b200.ops import RowLayout
from gslam_b200.optim import FusedAdam
from gslam_b200.synthetic import make_scene, target_images

import cpu_stubs as S

pytestmark = pytest.mark.gpu


def _ref_step(splats, **kw):
    adam = S.TorchRowsAdam(splats.params, [DEFAULT_LRS[n] for n in PARAM_NAMES])
    return MappingStep(splats, None, render_fn=S.oracle_render, loss_fn=S.oracle_loss, update_fn=adam.update,
                       collect_grads=S.collect_grads, **kw)


def _adam(splats, capturable=False):
    return FusedAdam([dict(params=[p], lr=DEFAULT_LRS[n]) for p, n in zip(splats.params, PARAM_NAMES)],
                     capturable=capturable)


@pytest.mark.parametrize("capturable", [False, True])
def test_mapping_step_matches_oracle_step(cuda, capturable):
    sc = make_scene(1500, 3, "tum", 0.2)
    gt = target_images(3, sc.height, sc.width)
    exposure = torch.tensor([[0.02, 0.01], [-0.01, 0.0], [0.0, -0.02]])
    ref = FlatSplats(sc.params())
    ref_step = _ref_step(ref, depth_tv_weight=1e-6)
    g = sc.to(cuda)
    ours = FlatSplats(g.params())
    our_step = MappingStep(ours, _adam(ours, capturable), depth_tv_weight=1e-6)
    lay = RowLayout(1)
    for it in range(2):
        r0 = ref_step(sc.viewmats, sc.Ks, gt, exposure, sc.width, sc.height, 3)
        r1 = our_step(g.viewmats, g.Ks, gt.to(cuda), exposure.to(cuda), sc.width, sc.height, 3)
        assert abs(float(r1.total) - float(r0.total)) < 1e-4 * abs(float(r0.total))
        assert abs(float(r1.photometric) - float(r0.photometric)) < 1e-4 * abs(float(r0.photometric))
        gr, go = ref_step._rows[:, :lay.vis], our_step._rows[:, :lay.vis].cpu()
        assert float((go - gr).norm() / gr.norm()) < 1e-3
        assert (r1.vis_count.cpu() != r0.vis_count).float().mean() < 1e-3
        assert r1.vis_count.dtype == torch.int32
    # Adam's first steps move every coordinate by ~lr regardless of the gradient's size, so compare
    # the parameter displacement, not the parameters
    p0 = FlatSplats(sc.params()).flat
    dr, do = ref.flat.detach() - p0, ours.flat.detach().cpu() - p0
    assert float((do - dr).norm() / dr.norm()) < 2e-2


def test_map_update_kernel_matches_the_torch_restatement(cuda):
    """gsb_map_update on given rows (one range, then two disjoint ranges as two ranks would run it)."""
    n = 10_007
    sc = make_scene(n, 1, "tum", 0.1)
    g = torch.Generator().manual_seed(0)
    lay = RowLayout(1)
    ref = FlatSplats(sc.params())
    ref_step = _ref_step(ref)
    ours = FlatSplats(sc.to(cuda).params())
    opt = _adam(ours)
    step = MappingStep(ours, opt)
    vis_ref = torch.zeros(n, dtype=torch.int32)
    vis = torch.zeros(n, dtype=torch.int32, device=cuda)
    for it in range(3):
        rows = torch.zeros(n, lay.stride)
        rows[:, :lay.vis] = torch.randn(n, lay.vis, generator=g) * 1e-2
        rows[:, lay.vis] = torch.randint(0, 4, (n,), generator=g).float()
        iso_ref = ref_step.update_fn(ref_step, rows, lay, 0, n, vis_ref)
        rows_d = rows.to(cuda)
        if it < 2:
            iso = opt.step_rows(ours.params, rows_d, lay, 0, n, vis, 2, 5e-4, 3, 0.995)
        else:       # two disjoint ranges, as two ranks run it (host step counter: one increment per call, undo one)
            a = opt.step_rows(ours.params, rows_d, lay, 0, 4000, vis, 2, 5e-4, 3, 0.995)
            for st in opt.state.values():
                st["step"] -= 1
            b = opt.step_rows(ours.params, rows_d, lay, 4000, n, vis, 2, 5e-4, 3, 0.995)
            iso = a + b
        assert abs(float(iso) - float(iso_ref)) < 1e-4 * abs(float(iso_ref))
        for p, q in zip(ours.params, ref.params):
            torch.testing.assert_close(p.detach().cpu(), q.detach(), rtol=2e-5, atol=1e-7)
        assert torch.equal(vis.cpu(), vis_ref)


def test_sh_mapping_step_matches_oracle_step(cuda):
    """BASELINE cfg4's parameterisation at a size the oracle handles: SH degree 3 coefficients [N,16,3]
    in the flat layout, rows of 64 floats, the wide-row branch of gsb_map_update."""
    n, C = 800, 2
    sc = make_scene(n, C, "tum", 0.2)
    g0 = torch.Generator().manual_seed(5)
    tensors = list(sc.params())
    sh = 0.1 * torch.randn(n, 16, 3, generator=g0)
    sh[:, 0] = tensors[4] * 0.5
    tensors[4] = sh
    gt = target_images(C, sc.height, sc.width)
    exposure = torch.zeros(C, 2)
    ref = FlatSplats(tensors)
    ref_step = _ref_step(ref, sh_degree=3)
    ours = FlatSplats([t.to(cuda) for t in tensors])
    our_step = MappingStep(ours, _adam(ours), sh_degree=3)
    lay = RowLayout(16)
    assert lay.stride == 64 and ours.widths[4] == 48
    g = sc.to(cuda)
    for it in range(2):
        r0 = ref_step(sc.viewmats, sc.Ks, gt, exposure, sc.width, sc.height, C)
        r1 = our_step(g.viewmats, g.Ks, gt.to(cuda), exposure.to(cuda), sc.width, sc.height, C)
        assert abs(float(r1.total) - float(r0.total)) < 1e-4 * abs(float(r0.total))
        gr, go = ref_step._rows[:, :lay.vis], our_step._rows[:, :lay.vis].cpu()
        assert float((go - gr).norm() / gr.norm()) < 1e-3
    p0 = FlatSplats(tensors).flat
    dr, do = ref.flat.detach() - p0, ours.flat.detach().cpu() - p0
    assert float((do - dr).norm() / dr.norm()) < 2e-2


def test_step_without_optimizer_then_apply_pending(cuda):
    sc = make_scene(2000, 2, "tum", 0.2).to(cuda)
    gt = target_images(2, sc.height, sc.width).to(cuda)
    ex = torch.zeros(2, 2, device=cuda)
    a, b = FlatSplats(sc.params()), FlatSplats(sc.params())
    sa, sb = MappingStep(a, _adam(a)), MappingStep(b, _adam(b))
    with __import__("gslam_b200").ops.deterministic_mode(True):
        sa(sc.viewmats, sc.Ks, gt, ex, sc.width, sc.height, 2)
        before = b.flat.clone()
        r = sb(sc.viewmats, sc.Ks, gt, ex, sc.width, sc.height, 2, step_optimizer=False)
        assert torch.equal(b.flat, before) and r.vis_count.dtype == torch.int32
        sb.apply_pending()
    assert torch.equal(a.flat, b.flat)


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
    ref_g, ref_loss = S.isotropic_grad(s, (radii > 0).sum(0) > 0, 5e-4)
    gpu_grad = grad0.clone().to(cuda)
    gpu_loss = add_isotropic_grad_(gpu_grad, s.to(cuda), vc, 5e-4)
    assert torch.allclose(gpu_grad.cpu(), grad0 + ref_g, rtol=1e-5, atol=1e-7)
    assert abs(float(gpu_loss) - float(ref_loss)) < 1e-4 * abs(float(ref_loss))
