"""Fused losses of the mapping and tracking steps, over libgsb.so.

* ``fused_ssim(img1, img2, padding='same', train=True)`` -- same call surface as the
  ``fused_ssim`` package the reference imports (gslam/backend.py:13, :303-307).
* ``photometric_ssim_loss`` -- the whole image-space part of one mapping iteration
  (exposure affine + active-NeRF NLL + 1 - SSIM, gslam/backend.py:273-283, :303-312)
  as one autograd node that hands ``v_render`` straight to the compositing backward.
* ``tracking_loss`` -- gslam/frontend.py:113-129 with the exposure affine of :632-636.
"""
from __future__ import annotations

import ctypes as ct
from typing import Optional

import torch
from torch import Tensor

from . import ops
from ._lib import call, lib, ptr, stream

PHOTO_MODES = {"mapping": 0, "active-nerf": 1, "l1": 2, "mse": 3}


def _strides(t: Tensor):
    return (ct.c_int64 * 4)(*t.stride())


def _nchw_like(t: Tensor) -> Tensor:
    """An uninitialised [N,C,H,W] tensor, channels-last in memory when `t` is."""
    N, C, H, W = t.shape
    if t.stride(1) < t.stride(3):
        return torch.empty(N, H, W, C, dtype=t.dtype, device=t.device).permute(0, 3, 1, 2)
    return torch.empty(N, C, H, W, dtype=t.dtype, device=t.device)


class _FusedSSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img1, img2, valid, train, det=False):
        assert img1.is_cuda and img1.dtype == torch.float32 and img1.dim() == 4 and img1.shape == img2.shape
        N, C, H, W = img1.shape
        if valid and (H <= 10 or W <= 10):
            raise ValueError("padding='valid' needs images larger than the 11x11 window")
        dev = img1.device
        ssim_sum = torch.zeros(1, dtype=torch.float32, device=dev)
        need = train and img1.requires_grad
        maps = torch.empty(3, N, C, H, W, dtype=torch.float32, device=dev) if need else None
        part = ops.det_scratch(lib().gsb_ssim_fwd_det_floats(N, C, H, W), dev) if det else None
        call("gsb_ssim_fwd", ptr(img1), ptr(img2), _strides(img1), _strides(img2), N, C, H, W, int(valid),
             ptr(ssim_sum), ptr(maps[0]) if need else None, ptr(maps[1]) if need else None,
             ptr(maps[2]) if need else None, ptr(part), stream())
        count = N * C * ((H - 10) * (W - 10) if valid else H * W)
        ctx.save_for_backward(img1, img2, maps)
        ctx.count = count
        return ssim_sum[0] / count

    @staticmethod
    def backward(ctx, g):
        img1, img2, maps = ctx.saved_tensors
        if maps is None:
            raise RuntimeError("fused_ssim(train=False) cannot be differentiated")
        N, C, H, W = img1.shape
        v = _nchw_like(img1)
        g = g.contiguous()
        call("gsb_ssim_bwd", ptr(img1), ptr(img2), _strides(img1), _strides(img2), N, C, H, W,
             ptr(maps[0]), ptr(maps[1]), ptr(maps[2]), ptr(g), 1.0 / ctx.count, ptr(v), _strides(v), 0,
             stream())
        return v, None, None, None, None


def fused_ssim(img1: Tensor, img2: Tensor, padding: str = "same", train: bool = True,
               deterministic: Optional[bool] = None) -> Tensor:
    """Mean SSIM of NCHW images (any strides); gradient to img1 only, like the original.
    deterministic (extension; None: ops.deterministic): fixed-order sum of the CTA partials."""
    assert padding in ("same", "valid"), padding
    det = ops.deterministic if deterministic is None else bool(deterministic)
    return _FusedSSIM.apply(img1, img2, padding == "valid", train, det)


class _PhotometricSSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, render, gt, exposure, ssim_weight, mode, unit_upstream=False, det=False, scale=1.0):
        ctx.set_materialize_grads(False)
        # render [C,H,W,5] = (r, g, b, depth, beta); gt [C,H,W,3]; exposure [C,2] or None
        assert render.is_cuda and render.is_contiguous() and render.shape[-1] == 5
        gt = gt.contiguous()
        C, H, W, _ = render.shape
        dev = render.device
        exposure_c = exposure.contiguous() if exposure is not None else None
        loss_sum = torch.zeros(3, dtype=torch.float32, device=dev)     # two sums + the block ticket
        loss_out = torch.empty(2, dtype=torch.float32, device=dev)     # (weighted, plain) photometric loss
        v_render = torch.empty_like(render)
        need_exp = exposure is not None and ctx.needs_input_grad[2]
        v_exp = torch.zeros(C, 2, dtype=torch.float32, device=dev) if need_exp else None
        part = ops.det_scratch(lib().gsb_photo_loss_det_floats(C, H, W), dev) if det else None
        # `scale` (a view-sharded step's share of the window) is folded into every coefficient: the value and
        # the gradient written here are already scaled, so no extra pass multiplies v_render later
        call("gsb_photo_loss", ptr(render), ptr(gt), ptr(exposure_c), mode, C, H, W, (1.0 - ssim_weight) * scale,
             ptr(loss_sum), ptr(loss_out), ptr(v_render), ptr(v_exp), ptr(part), stream())
        total, photo = loss_out[0], loss_out[1]
        ssim = None
        maps = img1 = img2 = None
        if ssim_weight != 0.0:
            img1 = render[..., :3].permute(0, 3, 1, 2)
            img2 = gt.permute(0, 3, 1, 2)
            ssim_sum = torch.zeros(1, dtype=torch.float32, device=dev)
            maps = torch.empty(3, C, 3, H, W, dtype=torch.float32, device=dev)
            spart = ops.det_scratch(lib().gsb_ssim_fwd_det_floats(C, 3, H, W), dev) if det else None
            call("gsb_ssim_fwd", ptr(img1), ptr(img2), _strides(img1), _strides(img2), C, 3, H, W, 1,
                 ptr(ssim_sum), ptr(maps[0]), ptr(maps[1]), ptr(maps[2]), ptr(spart), stream())
            ctx.count = C * 3 * (H - 10) * (W - 10)
            # total + w * (1 - ssim_sum / count) in two launches
            total = torch.sub(total, ssim_sum[0], alpha=scale * ssim_weight / ctx.count) + scale * ssim_weight
        ctx.save_for_backward(render, gt, maps, v_render, v_exp)
        ctx.ssim_weight = ssim_weight * scale
        ctx.unit_upstream = bool(unit_upstream) and ssim_weight == 0.0
        ctx.mark_non_differentiable(photo)
        return total, photo

    @staticmethod
    def backward(ctx, g, _g_photo):
        render, gt, maps, v_render, v_exp = ctx.saved_tensors
        C, H, W, _ = render.shape
        g = g.contiguous()
        if ctx.unit_upstream:
            # the caller calls backward() on this loss itself: the incoming gradient is exactly 1
            return v_render, None, v_exp, None, None, None, None, None
        # the saved gradient buffer is scaled and accumulated into IN PLACE below and returned as is (zero-copy);
        # a second backward through this node (retain_graph, checkpointing) would scale and accumulate twice
        if getattr(ctx, "consumed", False):
            raise RuntimeError("photometric_ssim_loss / tracking_loss: backward() ran twice through the same node; "
                               "its gradient buffer is single-use (recompute the loss instead of retain_graph=True)")
        ctx.consumed = True
        call("gsb_scale_by_scalar", ptr(v_render), v_render.numel(), ptr(g), stream())
        if maps is not None:
            img1 = render[..., :3].permute(0, 3, 1, 2)
            img2 = gt.permute(0, 3, 1, 2)
            v1 = v_render[..., :3].permute(0, 3, 1, 2)
            call("gsb_ssim_bwd", ptr(img1), ptr(img2), _strides(img1), _strides(img2), C, 3, H, W,
                 ptr(maps[0]), ptr(maps[1]), ptr(maps[2]), ptr(g), -ctx.ssim_weight / ctx.count, ptr(v1),
                 _strides(v1), 1, stream())
        return v_render, None, (v_exp * g if v_exp is not None else None), None, None, None, None, None


def photometric_ssim_loss(render: Tensor, gt: Tensor, exposure: Optional[Tensor] = None,
                          ssim_weight: float = 0.2, mode: str = "mapping",
                          deterministic: Optional[bool] = None, scale: float = 1.0):
    """(1 - w) * photometric + w * (1 - SSIM_valid)  ->  (total, photometric value).

    `render` is the [C,H,W,5] tensor behind RasterizationOutput (rgbs/depthmaps/betas are
    views of it; rasterization() attaches it as ``out.render``)."""
    det = ops.deterministic if deterministic is None else bool(deterministic)
    return _PhotometricSSIM.apply(render, gt, exposure, float(ssim_weight), PHOTO_MODES[mode], False, det, float(scale))


def tracking_loss(render: Tensor, gt: Tensor, exposure: Optional[Tensor] = None,
                  kind: str = "active-nerf", unit_upstream: bool = False,
                  deterministic: Optional[bool] = None) -> Tensor:
    """gslam/frontend.py:113-129 on the [1,H,W,5] render, exposure = (a, b) as [2] or [1,2].

    unit_upstream=True is the caller's promise that the loss reaches `backward()` with coefficient
    exactly 1 (`loss.backward()`, or `(loss + other).backward()`), as every tracking loop of the
    reference does: the backward then hands on the gradient the forward kernel already wrote instead
    of multiplying 1.5 M floats by 1.0 in one more pass.  A scaled use of such a loss is WRONG."""
    if exposure is not None and exposure.dim() == 1:
        exposure = exposure.view(1, 2)
    det = ops.deterministic if deterministic is None else bool(deterministic)
    total, _ = _PhotometricSSIM.apply(render, gt, exposure, 0.0, PHOTO_MODES[kind], unit_upstream, det)
    return total
