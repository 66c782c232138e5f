"""GPU parity of the fused SSIM / photometric losses and the fused Adam against the oracle."""
import math

import pytest
import torch

from gslam_b200.losses import fused_ssim, photometric_ssim_loss, tracking_loss
from gslam_b200.optim import FusedAdam
from oracle import torch_ref as R

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-12))


def _imgs(n, c, h, w, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(n, c, h, w, generator=g)
    b = (a + 0.2 * torch.randn(n, c, h, w, generator=g)).clamp(0, 1)
    return a, b


@pytest.mark.parametrize("padding", ["same", "valid"])
@pytest.mark.parametrize("shape", [(2, 3, 48, 64), (1, 3, 37, 53), (3, 1, 16, 16)])
def test_fused_ssim_matches_oracle(cuda, padding, shape):
    a, b = _imgs(*shape)
    ar = a.clone().requires_grad_(True)
    ref = R.fused_ssim(ar, b, padding=padding)
    ref.backward()
    ag = a.to(cuda).requires_grad_(True)
    out = fused_ssim(ag, b.to(cuda), padding=padding)
    out.backward()
    assert abs(out.item() - ref.item()) < 1e-5
    assert rel_err(ag.grad, ar.grad) < 1e-3   # gradient tolerance of north_star


def test_fused_ssim_channels_last_view_and_identity(cuda):
    a, b = _imgs(2, 3, 40, 56, seed=1)
    # NHWC memory viewed as NCHW, exactly what gslam passes (rgbs.permute(0, 3, 1, 2))
    nhwc = torch.zeros(2, 40, 56, 5, device=cuda)
    nhwc[..., :3] = a.permute(0, 2, 3, 1).to(cuda)
    nhwc.requires_grad_(True)
    out = fused_ssim(nhwc[..., :3].permute(0, 3, 1, 2), b.to(cuda), padding="valid")
    out.backward()
    ar = a.clone().requires_grad_(True)
    ref = R.fused_ssim(ar, b, padding="valid")
    ref.backward()
    assert abs(out.item() - ref.item()) < 1e-5
    assert rel_err(nhwc.grad[..., :3].permute(0, 3, 1, 2), ar.grad) < 1e-3
    assert (nhwc.grad[..., 3:] == 0).all()
    same = fused_ssim(a.to(cuda), a.to(cuda), padding="same", train=False)
    assert abs(same.item() - 1.0) < 1e-6


def _render(c, h, w, seed=0):
    g = torch.Generator().manual_seed(seed)
    render = torch.rand(c, h, w, 5, generator=g)
    render[..., 4] = 0.5 + 2.5 * render[..., 4]          # betas around e^1
    gt = torch.rand(c, h, w, 3, generator=g)
    exposure = 0.1 * torch.randn(c, 2, generator=g)
    return render, gt, exposure


def test_photometric_ssim_loss_matches_mapping_loss(cuda):
    """(1-w) * active-NeRF NLL + w * (1 - SSIM_valid), gslam/backend.py:273-283, :303-312."""
    render, gt, exposure = _render(2, 48, 64)
    rr = render.clone().requires_grad_(True)
    er = exposure.clone().requires_grad_(True)
    photo_ref = R.mapping_photometric_loss(R.exposure_affine(rr[..., :3], er), gt, rr[..., 4])
    ssim_ref = 1.0 - R.fused_ssim(rr[..., :3].permute(0, 3, 1, 2), gt.permute(0, 3, 1, 2), padding="valid")
    total_ref = 0.8 * photo_ref + 0.2 * ssim_ref
    (3.0 * total_ref).backward()
    rg = render.to(cuda).requires_grad_(True)
    eg = exposure.to(cuda).requires_grad_(True)
    total, photo = photometric_ssim_loss(rg, gt.to(cuda), eg, ssim_weight=0.2, mode="mapping")
    (3.0 * total).backward()
    assert abs(total.item() - total_ref.item()) < 1e-5 * max(1.0, abs(total_ref.item()))
    assert abs(photo.item() - photo_ref.item()) < 1e-5 * max(1.0, abs(photo_ref.item()))
    assert rel_err(rg.grad, rr.grad) < 1e-3
    assert rel_err(eg.grad, er.grad) < 1e-3


@pytest.mark.parametrize("kind", ["active-nerf", "l1", "mse"])
@pytest.mark.parametrize("unit", [False, True])
def test_tracking_loss_modes(cuda, kind, unit):
    render, gt, exposure = _render(1, 32, 48, seed=3)
    exposure = exposure[0]
    rr = render.clone().requires_grad_(True)
    er = exposure.clone().requires_grad_(True)
    img = rr[0, ..., :3] * er[0].exp() + er[1]
    err = img - gt[0]
    if kind == "active-nerf":
        ref = R.tracking_loss(gt[0], img, rr[0, ..., 4])
    elif kind == "l1":
        ref = err.abs().mean()
    else:
        ref = err.square().mean()
    ref.backward()
    rg = render.to(cuda).requires_grad_(True)
    eg = exposure.to(cuda).requires_grad_(True)
    # unit=True: the caller promises backward() starts at this loss, so the x1.0 pass is skipped
    out = tracking_loss(rg, gt.to(cuda), eg, kind=kind, unit_upstream=unit)
    out.backward()
    assert abs(out.item() - ref.item()) < 1e-5 * max(1.0, abs(ref.item()))
    assert rel_err(rg.grad, rr.grad) < 1e-3
    assert rel_err(eg.grad, er.grad) < 1e-3


@pytest.mark.parametrize("capturable", [False, True])
def test_fused_adam_matches_torch_adam(cuda, capturable):
    """Six tensors with the reference's learning rates (gslam/backend.py:53-58), 5 steps, one launch each;
    capturable=True keeps the step count on the device (CUDA-graph replay)."""
    g = torch.Generator().manual_seed(0)
    n = 1000
    shapes = [(n, 3), (n, 4), (n, 3), (n,), (n, 3), (n,)]
    lrs = [1.6e-3, 5e-3, 5e-3, 2.5e-2, 1e-2, 2.5e-3]
    init = [torch.randn(s, generator=g) for s in shapes]
    ref_p = [p.clone().requires_grad_(True) for p in init]
    ref_opts = [torch.optim.Adam([p], lr=lr) for p, lr in zip(ref_p, lrs)]
    our_p = [p.clone().to(cuda).requires_grad_(True) for p in init]
    ours = FusedAdam([dict(params=[p], lr=lr) for p, lr in zip(our_p, lrs)], capturable=capturable)
    vis = torch.randint(0, 4, (n,), generator=g, dtype=torch.int32)
    for it in range(5):
        grads = [torch.randn(s, generator=g) for s in shapes]
        for p, q, gr in zip(ref_p, our_p, grads):
            p.grad = gr.clone()
            q.grad = gr.to(cuda)
        for o in ref_opts:
            o.step()
        with torch.no_grad():
            ref_p[3].data[vis > 1] *= 0.995
        ours.step(decay_param=our_p[3], vis_count=vis.to(cuda), decay=0.995)
    for p, q in zip(ref_p, our_p):
        torch.testing.assert_close(q.detach().cpu(), p.detach(), rtol=1e-5, atol=1e-6)
    st = ours.state[our_p[0]]
    assert set(st) == {"step", "exp_avg", "exp_avg_sq"} and st["exp_avg"].shape == (n, 3)
    assert int(st["step"]) == 5
    torch.testing.assert_close(st["exp_avg"].cpu(), ref_opts[0].state[ref_p[0]]["exp_avg"], rtol=1e-5, atol=1e-7)

