"""View-sharded mapping step: the body of Backend.optimize_map (reference
gslam/backend.py:260-359) with the keyframe window split across ranks.

The reference is single-GPU.  Each view's projection / binning / compositing /
image loss is independent given the (replicated) Gaussians, so rank r of p
renders views window[r::p] and ONE all-reduce of the flat [15 N] gradient buffer
(means | quats | scales | opacities | colors | log_uncertainties) per step makes
every rank apply the identical Adam update; parameters never need a broadcast
after step 0.  Pose / exposure gradients stay on the rank that owns the view.

Host-side pieces (sharding, loss scaling, flat-gradient reduction, the analytic
isotropic-regulariser gradient) are plain functions so they can be tested on the
CPU with the gloo backend; the renderer is `gslam_b200.rasterization`.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist
from torch import Tensor

PARAM_NAMES = ("means", "quats", "scales", "opacities", "colors", "log_uncertainties")
PARAM_WIDTH = (3, 4, 3, 1, 3, 1)          # 15 floats per Gaussian
# learning rates of reference MapConfig (gslam/backend.py:53-58)
DEFAULT_LRS = dict(means=1.6e-3, quats=5e-3, scales=5e-3, opacities=2.5e-2, colors=1e-2,
                   log_uncertainties=2.5e-3)


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Views of the window rendered by `rank`: window[rank::world]."""
    return list(range(rank, n_views, world))


class FlatSplats:
    """The six per-splat tensors as views of ONE flat [15 N] buffer (same for the gradients)."""

    def __init__(self, tensors: Sequence[Tensor]):
        n = tensors[0].shape[0]
        dev, dt = tensors[0].device, tensors[0].dtype
        self.n = n
        self.flat = torch.empty(15 * n, device=dev, dtype=dt)
        self.params: List[Tensor] = []
        o = 0
        for t, w in zip(tensors, PARAM_WIDTH):
            v = self.flat[o:o + w * n].view((n, w) if w > 1 else (n,))
            v.copy_(t.detach().reshape(v.shape))
            self.params.append(v.requires_grad_(True))
            o += w * n

    def flat_grad(self, tail: int = 0) -> Tensor:
        """The gradient as ONE buffer [15 N + tail]: the six gradients in parameter order, then `tail`
        scratch floats (uninitialised) that ride along in the same all-reduce.  Zero-copy when the
        grads are still the six consecutive slices of project_bwd's flat buffer and that buffer was
        allocated with room for the tail (ops.grad_tail_floats); autograd hands the grads over
        detached, so this is decided from the storage, not from `_base`."""
        grads = [p.grad for p in self.params]
        g0 = grads[0]
        total = 15 * self.n + tail
        if g0 is not None and all(g is not None and g.is_contiguous() for g in grads):
            st, o, ok = g0.untyped_storage(), g0.storage_offset(), True
            expect = o
            for g, w in zip(grads, PARAM_WIDTH):
                ok = ok and g.untyped_storage().data_ptr() == st.data_ptr() and g.storage_offset() == expect
                expect += w * self.n
            if ok and st.nbytes() >= (o + total) * g0.element_size():
                return torch.empty(0, dtype=g0.dtype, device=g0.device).set_(st, o, (total,), (1,))
        flat = torch.empty(total, dtype=self.flat.dtype, device=self.flat.device)
        o = 0
        for p, g, w in zip(self.params, grads, PARAM_WIDTH):
            dst = flat[o:o + w * self.n].view(p.shape)
            if g is None:
                dst.zero_()
            else:
                dst.copy_(g)
            p.grad = dst
            o += w * self.n
        return flat


@torch.no_grad()
def visibility_count(radii: Tensor) -> Tensor:
    """(radii > 0).sum(dim=0) as int32 [N] (gslam/backend.py:287, :357): one launch on CUDA."""
    if radii.is_cuda:
        from ._lib import call, ptr, stream
        C, N = radii.shape
        out = torch.empty(N, dtype=torch.int32, device=radii.device)
        call("gsb_vis_count", ptr(radii.contiguous()), C, N, ptr(out), stream())
        return out
    return (radii > 0).sum(dim=0).to(torch.int32)


@torch.no_grad()
def add_isotropic_grad_(scales_grad: Tensor, log_scales: Tensor, visible: Tensor, weight: float):
    """d/d log_scales of  weight * sum_visible |exp(s) - exp(mean_axis s).detach()|
    (gslam/backend.py:286-296), added in place; returns the loss value.  `visible` is the int32
    visibility count (or a bool mask).  CUDA tensors take one kernel (gsb_isotropic_reg); the torch
    expression below serves the CPU host-logic tests of the mapping step (gloo, stub renderer)."""
    if scales_grad.is_cuda and visible.dtype == torch.int32 and scales_grad.is_contiguous():
        from ._lib import call, ptr, stream
        from . import ops
        loss = torch.zeros(1, dtype=torch.float32, device=scales_grad.device)
        n = log_scales.shape[0]
        part = ops.det_scratch((n + 255) // 256, scales_grad.device) if ops.deterministic else None
        call("gsb_isotropic_reg", ptr(log_scales.detach().contiguous()), ptr(visible), n,
             float(weight), ptr(scales_grad), ptr(loss), ptr(part), stream())
        return loss[0]
    visible = visible > 0 if visible.dtype != torch.bool else visible
    s = log_scales.detach()
    e = s.exp()
    m = s.mean(dim=1, keepdim=True).exp()
    vis = visible[:, None].to(s.dtype)
    scales_grad.add_(weight * torch.sign(e - m) * e * vis)
    return ((e - m).abs() * vis).sum()


def edge_aware_tv(depth: Tensor, rgb: Tensor, mask: Tensor) -> Tensor:
    """Edge-aware total variation of the depth map (reference gslam/utils.py:136-161):
    |d depth| * exp(-mean_ch |d rgb|) summed over the masked horizontal and vertical neighbours."""
    ddx = (depth[..., :, :-1] - depth[..., :, 1:]).abs()
    ddy = (depth[..., :-1, :] - depth[..., 1:, :]).abs()
    wx = torch.exp(-(rgb[..., :, :-1, :] - rgb[..., :, 1:, :]).abs().mean(-1))
    wy = torch.exp(-(rgb[..., :-1, :, :] - rgb[..., 1:, :, :]).abs().mean(-1))
    return (ddx * wx * mask[..., :, :-1]).sum() + (ddy * wy * mask[..., :-1, :]).sum()


LOCAL = "local"      # MappingStep(group=LOCAL): this rank maps on its own even inside a multi-rank job


def allreduce_sum_(t: Tensor, group=None) -> Tensor:
    if isinstance(group, str) and group == LOCAL:
        return t
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


@dataclass
class MappingStepResult:
    total: Tensor
    photometric: Tensor
    radii: Tensor
    n_touched: Tensor
    vis_count: Tensor
    outputs: Optional[object] = None     # this rank's RasterizationOutput (est. depths, means2d.grad, ...)


class MappingStep:
    """fwd + losses + bwd + gradient all-reduce + Adam + opacity decay for this rank's views."""

    def __init__(self, splats: FlatSplats, optimizer, render_fn: Optional[Callable] = None,
                 loss_fn: Optional[Callable] = None, ssim_weight: float = 0.2,
                 isotropic_weight: float = 5e-4, opacity_decay: float = 0.995, depth_tv_weight: float = 0.0,
                 group=None):
        if render_fn is None:
            from .rasterization import rasterization as render_fn
        if loss_fn is None:
            from .losses import photometric_ssim_loss as loss_fn
        self.splats, self.opt = splats, optimizer
        self.render_fn, self.loss_fn = render_fn, loss_fn
        self.ssim_weight, self.iso_w, self.decay, self.group = ssim_weight, isotropic_weight, opacity_decay, group
        self.tv_w = depth_tv_weight
        self.retain_means2d_grad = False      # out.means2d.retain_grad() of gslam/backend.py:326 (densification)

    def __call__(self, viewmats: Tensor, Ks: Tensor, gt_imgs: Tensor, exposure: Optional[Tensor],
                 width: int, height: int, n_views_total: int, background: Optional[Tensor] = None,
                 step_optimizer: bool = True, gt_depths: Optional[Tensor] = None) -> MappingStepResult:
        P = self.splats.params
        for p in P:
            p.grad = None
        c_local = viewmats.shape[0]
        bg = background if background is not None else torch.zeros(c_local, 3, device=viewmats.device)
        out = self.render_fn(P[0], P[1], P[2], P[3], P[4], viewmats, Ks, width, height, packed=False,
                             render_mode="RGB+D", log_uncertainties=P[5], backgrounds=bg)
        # the image losses are means over ALL views of the window: weight this rank's share
        total, photo = self.loss_fn(out.render, gt_imgs, exposure, self.ssim_weight, "mapping")
        share = c_local / float(n_views_total)
        loss = total * share
        if self.tv_w:
            # a sum (not a mean) over this rank's views: no share factor (gslam/backend.py:297-318)
            loss = loss + self.tv_w * edge_aware_tv(out.depthmaps, out.rgbs, out.alphas[..., 0] > 0.4)
        if gt_depths is not None:
            # use_gt_depths (gslam/backend.py:320-323): 0.1 * mean |depth - gt| over gt > 0, a mean over
            # the window's valid pixels: this rank's share is its valid-pixel fraction of the window total
            valid = (gt_depths > 0).to(out.depthmaps.dtype)
            n_valid = allreduce_sum_(valid.sum().view(1).clone(), self.group).clamp_min(1.0)
            loss = loss + 0.1 * ((out.depthmaps - gt_depths).abs() * valid).sum() / n_valid[0]
        from . import ops
        n = self.splats.n
        ops.grad_tail_floats = n + 2          # room behind the flat gradient for counts and loss scalars
        if self.retain_means2d_grad:
            out.means2d.retain_grad()
        try:
            loss.backward()
        finally:
            ops.grad_tail_floats = 0
        # ONE collective per step: [15 N gradient | N visibility counts | total, photometric]
        buf = self.splats.flat_grad(tail=n + 2)
        t_vis, t_stats = buf[15 * n:16 * n], buf[16 * n:]
        t_vis.copy_(visibility_count(out.radii))            # exact in float32 (counts <= #views)
        t_stats.copy_(torch.stack([total.detach(), photo.detach()]) * share)
        allreduce_sum_(buf, self.group)
        vis_count = t_vis.to(torch.int32)
        stats = t_stats
        # isotropic regulariser on the union visibility, identical on every rank, after the reduce
        iso = add_isotropic_grad_(P[2].grad, P[2], vis_count, self.iso_w)
        if step_optimizer:
            self.opt.step(decay_param=P[3], vis_count=vis_count, decay=self.decay)
        return MappingStepResult(stats[0] + self.iso_w * iso, stats[1], out.radii, out.n_touched, vis_count, out)
