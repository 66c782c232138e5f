"""GPU parity of the pose chain (rows a9/a10 of SURVEY 8a) against the oracle.

The oracle's pose_zhou / rotation_6d_to_matrix restate gslam/primitives.py:15-36, 82-92 in torch on
the CPU; gradients there come from autograd, here from the hand-written backward kernel.
"""
import pytest
import torch

from gslam_b200.map import GaussianSplattingData
from gslam_b200.primitives import Camera, PoseZhou, pose_batch, pose_chain, rotation_6d_to_matrix
from gslam_b200.rasterization import rasterization
from gslam_b200.synthetic import make_scene
from oracle import torch_ref as R

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-12))


def _random_rt(C, gen):
    q = torch.randn(C, 6, generator=gen)
    rot = torch.stack([R.rotation_6d_to_matrix(x) for x in q])
    Rt = torch.eye(4).repeat(C, 1, 1)
    Rt[:, :3, :3] = rot
    Rt[:, :3, 3] = torch.randn(C, 3, generator=gen)
    return Rt


@pytest.mark.parametrize("C", [1, 8, 100])
@pytest.mark.parametrize("scale", [0.0, 0.05, 0.7])
def test_pose_chain_forward_backward(cuda, C, scale):
    gen = torch.Generator().manual_seed(C * 7 + int(scale * 100))
    Rt = _random_rt(C, gen)
    dR0, dt0 = scale * torch.randn(C, 6, generator=gen), scale * torch.randn(C, 3, generator=gen)
    w = torch.randn(C, 4, 4, generator=gen)
    dR_r, dt_r = dR0.clone().requires_grad_(True), dt0.clone().requires_grad_(True)
    ref = torch.stack([R.pose_zhou(Rt[c], dR_r[c], dt_r[c]) for c in range(C)])
    (ref * w).sum().backward()
    dR_g, dt_g = dR0.to(cuda).requires_grad_(True), dt0.to(cuda).requires_grad_(True)
    out = pose_chain(Rt.to(cuda), dR_g, dt_g)
    (out * w.to(cuda)).sum().backward()
    assert torch.allclose(out.cpu(), ref.detach(), rtol=1e-5, atol=1e-6)
    assert rel_err(dR_g.grad, dR_r.grad) < 1e-4      # north_star: gradients 1e-3 relative
    assert rel_err(dt_g.grad, dt_r.grad) < 1e-5
    # a rotation comes out: orthonormal, det +1
    rot = out[:, :3, :3].cpu().double()
    base = Rt[:, :3, :3].double()
    delta = base.transpose(1, 2) @ rot
    assert torch.allclose(delta @ delta.transpose(1, 2), torch.eye(3, dtype=torch.float64).expand(C, 3, 3), atol=1e-5)
    assert torch.allclose(torch.linalg.det(delta), torch.ones(C, dtype=torch.float64), atol=1e-5)


def test_rotation_6d_to_matrix_matches_oracle(cuda):
    d6 = torch.randn(5, 7, 6, generator=torch.Generator().manual_seed(3))
    ref = R.rotation_6d_to_matrix(d6)
    out = rotation_6d_to_matrix(d6.to(cuda))
    assert out.shape == (5, 7, 3, 3)
    assert torch.allclose(out.cpu(), ref, rtol=1e-5, atol=1e-6)


def test_pose_module_mirrors_reference_fields(cuda):
    Rt = _random_rt(1, torch.Generator().manual_seed(1))[0]
    pose = PoseZhou(Rt.clone()).to(cuda)
    assert set(dict(pose.named_parameters())) == {"dt", "dR"}
    assert set(dict(pose.named_buffers())) == {"Rt", "id", "eye4_row4"}
    assert torch.allclose(pose().cpu(), Rt, atol=1e-6)          # zero delta = the base pose
    frozen = PoseZhou(Rt.clone(), is_learnable=False).to(cuda)
    assert frozen() is frozen.Rt
    with torch.no_grad():
        pose.dR.copy_(torch.tensor([0.01, -0.02, 0.03, 0.02, 0.01, -0.01]))
        pose.dt.copy_(torch.tensor([0.1, -0.2, 0.05]))
    ref = R.pose_zhou(Rt, pose.dR.detach().cpu(), pose.dt.detach().cpu())
    assert torch.allclose(pose().cpu(), ref, rtol=1e-5, atol=1e-6)
    batch = pose_batch([pose, pose])
    assert batch.shape == (2, 4, 4) and torch.equal(batch[0], batch[1])
    mixed = pose_batch([pose, frozen])
    assert torch.allclose(mixed[1].cpu(), Rt, atol=1e-6)


def test_map_forward_routes_gradients_to_poses(cuda):
    """GaussianSplattingData.forward (gslam/map.py:78-103): pose parameters receive the viewmat gradient."""
    g = make_scene(3000, 2, "tum", 0.5).to(cuda)
    p = g.params()
    ages = torch.zeros(p[0].shape[0], dtype=torch.long, device=cuda)
    m = GaussianSplattingData(p[0].clone(), p[1].clone(), p[2].clone(), p[3].clone(), p[4].clone(), p[5].clone(), ages)
    cams = [Camera(g.Ks[c].clone(), g.height, g.width) for c in range(2)]
    poses = [PoseZhou(g.viewmats[c].clone()).to(cuda) for c in range(2)]
    with torch.no_grad():
        poses[1].dR.add_(0.01)
        poses[1].dt.add_(0.02)
    out = m(cams, poses, render_depth=True)
    w = torch.randn(out.render.shape, generator=torch.Generator().manual_seed(0)).to(cuda)
    (out.render * w).sum().backward()
    # the same render with explicit viewmats as leaves: chain rule through the oracle's pose map
    vm = torch.stack([q() for q in poses]).detach().requires_grad_(True)
    out2 = rasterization(m.means, m.quats, m.scales, m.opacities, m.colors, vm, g.Ks, g.width, g.height,
                         packed=False, render_mode="RGB+D", log_uncertainties=m.log_uncertainties,
                         backgrounds=torch.zeros(2, 3, device=cuda))
    assert torch.equal(out.render, out2.render)
    (out2.render * w).sum().backward()
    for c in range(2):
        dR = poses[c].dR.detach().cpu().requires_grad_(True)
        dt = poses[c].dt.detach().cpu().requires_grad_(True)
        (R.pose_zhou(poses[c].Rt.cpu(), dR, dt) * vm.grad[c].cpu()).sum().backward()
        assert rel_err(poses[c].dR.grad, dR.grad) < 1e-3
        assert rel_err(poses[c].dt.grad, dt.grad) < 1e-3
    out_r = m.render(cams, [q().detach() for q in poses])
    assert torch.equal(out_r.render, out.render)
    assert m.mask(torch.arange(10, device=cuda)).means.shape == (10, 3)
    assert not m.no_grad_clone().means.requires_grad and set(m.as_dict()) == set(m._per_splat_params)
