"""View-sharded mapping step: the body of Backend.optimize_map (reference
gslam/backend.py:260-359) with the keyframe window split across ranks.

The reference is single-GPU.  Each view's projection / binning / compositing / image loss is
independent given the (replicated) Gaussians, so rank r of p renders views window[r::p].  Per step:

  1. backward writes the gradient as row-major records ``rows[N_pad, stride]`` -- a Gaussian's six
     gradients and its visibility count in ONE contiguous row (`ops.RowLayout`, `gsb_project_bwd`);
  2. ``reduce_scatter`` of the rows: rank r receives the summed rows of ITS Gaussian range [n0, n1);
  3. ONE kernel (`gsb_map_update`) on that range: union visibility -> isotropic-regulariser gradient
     -> Adam on all six parameters -> opacity decay;
  4. ``all_gather`` of the updated parameter rows (six blocks + the visibility counts + three loss
     scalars, one coalesced NCCL group).

An all-reduce IS a reduce-scatter plus an all-gather, so (2) + (4) move the bytes round 1's single
all-reduce of the flat gradient moved, while the O(N) tail of the step -- visibility count,
regulariser, Adam, decay: three kernels over all N on every rank before -- runs once over N / p
Gaussians.  With one rank there is no collective and (3) runs over [0, N).  Pose / exposure
gradients stay on the rank that owns the view.

The collectives go through `torch.distributed` (NCCL over NVLink on the GPUs; the host logic is
exercised with gloo in tests/test_mapping_cpu.py, where the renderer, the loss and the update come
from the oracle).  There is no CPU branch in this module: without an injected `update_fn` /
`collect_grads` everything below runs the CUDA kernels.
"""
from __future__ import annotations

import contextlib
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist
from torch import Tensor

from . import ops
from .ops import RowLayout

PARAM_NAMES = ("means", "quats", "scales", "opacities", "colors", "log_uncertainties")
PARAM_WIDTH = (3, 4, 3, 1, 3, 1)          # floats per Gaussian without SH (15); colours are 3 K with SH
# learning rates of reference MapConfig (gslam/backend.py:53-58)
DEFAULT_LRS = dict(means=1.6e-3, quats=5e-3, scales=5e-3, opacities=2.5e-2, colors=1e-2,
                   log_uncertainties=2.5e-3)


def symmetric_zeros(numel: int, dtype, device, group):
    """A zeroed torch symmetric-memory tensor and its rendezvous handle (collective over `group`: every
    rank calls it with the same size, in the same order).  The handle carries the peers' mapped addresses,
    the NVLS multicast address (0 if the fabric has none) and a device-side barrier."""
    import torch.distributed._symmetric_memory as symm_mem
    t = symm_mem.empty(max(int(numel), 1), dtype=dtype, device=device)
    h = symm_mem.rendezvous(t, group=group if group is not None else dist.group.WORLD)
    t.zero_()
    return t[:numel], h


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Views of the window rendered by `rank`: window[rank::world]."""
    return list(range(rank, n_views, world))


def param_widths(K: int = 1):
    return (3, 4, 3, 1, 3 * K, 1)


def _round_up(n: int, a: int) -> int:
    return -(-n // a) * a


class FlatSplats:
    """The six per-splat tensors as views of ONE flat buffer: six blocks of `cap` rows each
    (means | quats | scales | opacities | colors | log_uncertainties), `cap` = n rounded up to `align`
    (the number of ranks of a view-sharded step: every block then splits into equal row ranges for the
    all-gather of the updated parameters); a parameter is the first n rows of its block.  Colours are
    [N,3] (the reference) or SH coefficients [N,K,3] (extension, BASELINE cfg4)."""

    def __init__(self, tensors: Sequence[Tensor], align: int = 1, symmetric=None):
        """symmetric: a process group -- the flat buffer is then a torch symmetric-memory allocation shared with
        the group's ranks (same size everywhere; `self.symm` is its handle), which lets the fused update
        kernel multicast-store the updated parameters into every GPU's copy (MappingStep, NVLS path)."""
        n = tensors[0].shape[0]
        dev, dt = tensors[0].device, tensors[0].dtype
        colors = tensors[4]
        self.K = 1 if colors.dim() == 2 else int(colors.shape[1])
        self.widths = param_widths(self.K)
        self.align = max(int(align), 1)
        self.n = n
        self.cap = _round_up(n, self.align)
        self.symm = None
        if symmetric is not None:
            self.flat, self.symm = symmetric_zeros(sum(self.widths) * self.cap, dt, dev, symmetric)
        else:
            self.flat = torch.zeros(sum(self.widths) * self.cap, device=dev, dtype=dt)
        self.params: List[Tensor] = []
        for k, v in enumerate(self.views(self.flat, n, self.cap, self.K)):
            v.copy_(tensors[k].detach().reshape(v.shape))
            self.params.append(v.requires_grad_(True))

    @staticmethod
    def views(flat: Tensor, n: int, cap: int, K: int = 1) -> List[Tensor]:
        """The six [n, ...] parameter views of a flat buffer laid out as blocks of `cap` rows."""
        out, o = [], 0
        for k, w in enumerate(param_widths(K)):
            v = flat[o:o + w * n]
            if k == 4 and K > 1:
                v = v.view(n, K, 3)
            else:
                v = v.view((n, w) if w > 1 else (n,))
            out.append(v)
            o += w * cap
        return out

    def blocks(self) -> List[Tensor]:
        """The six blocks at full capacity, flat ([w * cap] each): what the all-gather moves."""
        out, o = [], 0
        for w in self.widths:
            out.append(self.flat[o:o + w * self.cap])
            o += w * self.cap
        return out


@torch.no_grad()
def visibility_count(radii: Tensor) -> Tensor:
    """(radii > 0).sum(dim=0) as int32 [N] (gslam/backend.py:287, :357): one launch."""
    from ._lib import call, ptr, stream
    C, N = radii.shape
    out = torch.empty(N, dtype=torch.int32, device=radii.device)
    call("gsb_vis_count", ptr(radii.contiguous()), C, N, ptr(out), stream())
    return out


@torch.no_grad()
def add_isotropic_grad_(scales_grad: Tensor, log_scales: Tensor, visible: Tensor, weight: float):
    """d/d log_scales of  weight * sum_visible |exp(s) - exp(mean_axis s).detach()|
    (gslam/backend.py:286-296), added in place; returns the (unweighted) loss value.  `visible` is the
    int32 visibility count.  One kernel (gsb_isotropic_reg); the mapping step itself has this fused
    into gsb_map_update."""
    from ._lib import call, ptr, stream
    if visible.dtype != torch.int32:
        visible = visible.to(torch.int32)
    loss = torch.zeros(1, dtype=torch.float32, device=scales_grad.device)
    n = log_scales.shape[0]
    part = ops.det_scratch((n + 255) // 256, scales_grad.device) if ops.deterministic else None
    assert scales_grad.is_contiguous()
    call("gsb_isotropic_reg", ptr(log_scales.detach().contiguous()), ptr(visible.contiguous()), n,
         float(weight), ptr(scales_grad), ptr(loss), ptr(part), stream())
    return loss[0]


def edge_aware_tv(depth: Tensor, rgb: Tensor, mask: Tensor) -> Tensor:
    """Edge-aware total variation of the depth map (reference gslam/utils.py:136-161):
    |d depth| * exp(-mean_ch |d rgb|) summed over the masked horizontal and vertical neighbours."""
    ddx = (depth[..., :, :-1] - depth[..., :, 1:]).abs()
    ddy = (depth[..., :-1, :] - depth[..., 1:, :]).abs()
    wx = torch.exp(-(rgb[..., :, :-1, :] - rgb[..., :, 1:, :]).abs().mean(-1))
    wy = torch.exp(-(rgb[..., :-1, :, :] - rgb[..., 1:, :, :]).abs().mean(-1))
    return (ddx * wx * mask[..., :, :-1]).sum() + (ddy * wy * mask[..., :-1, :]).sum()


LOCAL = "local"      # MappingStep(group=LOCAL): this rank maps on its own even inside a multi-rank job


def _world(group):
    """(rank, world size) of `group`; (0, 1) for LOCAL or outside torch.distributed."""
    if isinstance(group, str) and group == LOCAL:
        return 0, 1
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def allreduce_sum_(t: Tensor, group=None) -> Tensor:
    if _world(group)[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def allreduce_max_(t: Tensor, group=None) -> Tensor:
    if _world(group)[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return t


@torch.no_grad()
def gather_adam_state(splats: "FlatSplats", optimizer, group=None) -> None:
    """Make every rank's Adam moments complete.  A view-sharded step updates the moments of Gaussian g only on the
    rank whose range [r * cap / p, (r + 1) * cap / p) holds g; everywhere else they are stale.  That is fine while
    the ranges stand still, but anything that changes N or the capacity (insertion, densification, pruning: the
    compaction moves every later Gaussian down, the new capacity moves every boundary) hands Gaussians to ranks
    that never updated their moments -- on 8 ranks with a few thousand splats most of the map after a few
    keyframes.  So before the Gaussians are re-partitioned every rank collects all moments by the CURRENT ranges:
    one all-gather of [cap / p, 2 * sum(widths)] floats.  After it any partition is valid."""
    rank, world = _world(group)
    if world == 1 or optimizer is None or splats.n == 0:
        return
    n, per = splats.n, splats.cap // world
    tensors = []
    for prm, w in zip(splats.params, splats.widths):
        st = optimizer.state.get(prm)
        if st and "exp_avg" in st:
            tensors += [(st["exp_avg"], w), (st["exp_avg_sq"], w)]
    if not tensors:
        return
    dev = splats.flat.device
    width = sum(w for _, w in tensors)
    send = torch.zeros(per, width, dtype=torch.float32, device=dev)
    lo, hi = min(rank * per, n), min((rank + 1) * per, n)
    col = 0
    for t, w in tensors:
        if hi > lo:
            send[:hi - lo, col:col + w] = t.reshape(n, w)[lo:hi]
        col += w
    recv = torch.empty(world * per, width, dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
    col = 0
    for t, w in tensors:
        t.reshape(n, w).copy_(recv[:n, col:col + w])
        col += w


def _coalesced(group, device):
    """One NCCL group call for several collectives (a capability of the backend, not a fallback: gloo has
    no coalescing and simply issues them one by one)."""
    if dist.get_backend(group) == "nccl":
        return dist._coalescing_manager(group=group, device=device, async_ops=False)
    return contextlib.nullcontext()


@dataclass
class MappingStepResult:
    total: Tensor
    photometric: Tensor
    radii: Tensor
    n_touched: Tensor
    vis_count: Tensor
    outputs: Optional[object] = None     # this rank's RasterizationOutput (est. depths, means2d.grad, ...)


def fused_update(step: "MappingStep", rows: Tensor, lay: RowLayout, n0: int, n1: int, vis_out: Tensor) -> Tensor:
    """The default `update_fn`: gsb_map_update through the optimizer that owns the Adam state.
    -> the rank's partial sum of the (unweighted) isotropic loss, a [1] tensor."""
    return step.opt.step_rows(step.splats.params, rows, lay, n0, n1, vis_out, scales_index=2,
                              iso_weight=step.iso_w, decay_index=3, decay=step.decay)


class MappingStep:
    """fwd + losses + bwd + reduce-scatter + fused (regulariser, Adam, decay) on this rank's Gaussian
    range + all-gather, for this rank's views.

    `render_fn`, `loss_fn`: default the CUDA rasterizer and the fused image losses.  `update_fn(step, rows,
    layout, n0, n1, vis_out)` and `collect_grads(params, radii, rows, layout)` exist so that the HOST logic
    (sharding, ranges, collectives, loss shares) can be tested without a GPU: the tests inject oracle-based
    stand-ins; the defaults are the kernels and `collect_grads=None` means "the backward pass wrote the rows
    itself" (checked)."""

    def __init__(self, splats: FlatSplats, optimizer, render_fn: Optional[Callable] = None,
                 loss_fn: Optional[Callable] = None, ssim_weight: float = 0.2,
                 isotropic_weight: float = 5e-4, opacity_decay: float = 0.995, depth_tv_weight: float = 0.0,
                 group=None, sh_degree: Optional[int] = None, update_fn: Optional[Callable] = None,
                 collect_grads: Optional[Callable] = None, comm: str = "auto"):
        """comm (more than one rank): "nccl" = reduce-scatter / all-gather collectives around the update
        kernel; "nvls" = ONE kernel that does all three over NVLink multicast (needs FlatSplats(symmetric=group)
        and a fabric with multicast: torch symmetric memory's multicast_ptr); "auto" = nvls when the splats
        are symmetric and multicast is there, else nccl."""
        if render_fn is None:
            from .rasterization import rasterization as render_fn
        self._fused_loss = loss_fn is None
        if loss_fn is None:
            from .losses import photometric_ssim_loss as loss_fn
        self.splats, self.opt = splats, optimizer
        self.render_fn, self.loss_fn = render_fn, loss_fn
        self.update_fn = update_fn or fused_update
        self.collect_grads = collect_grads
        self.ssim_weight, self.iso_w, self.decay, self.group = ssim_weight, isotropic_weight, opacity_decay, group
        self.tv_w = depth_tv_weight
        self.sh_degree = sh_degree
        self.retain_means2d_grad = False      # out.means2d.retain_grad() of gslam/backend.py:326 (densification)
        if comm not in ("auto", "nccl", "nvls"):
            raise ValueError(comm)
        self.comm = comm
        self.phase_events: Optional[list] = None   # bench.py: [(name, start, end)] CUDA events of an eager step
        self._buf_key = None
        self._pending = None

    # ---- buffers that live across steps (a CUDA graph of the step bakes their addresses in) -------------
    def _buffers(self, lay: RowLayout, world: int):
        sp = self.splats
        dev = sp.flat.device
        key = (sp.cap, lay.stride, world, dev)
        if key != self._buf_key:
            self._nvls = None
            use_nvls = world > 1 and self.comm != "nccl" and sp.symm is not None and \
                bool(getattr(sp.symm, "multicast_ptr", 0)) and self.update_fn is fused_update
            if self.comm == "nvls" and world > 1 and not use_nvls:
                raise RuntimeError("MappingStep(comm='nvls') needs FlatSplats(symmetric=group) on a fabric with "
                                   "NVLink multicast and the default update_fn")
            if use_nvls:
                rows, h_rows = symmetric_zeros(sp.cap * lay.stride, torch.float32, dev, self.group)
                vis, h_vis = symmetric_zeros(sp.cap, torch.int32, dev, self.group)
                stats, h_stats = symmetric_zeros(world * 4, torch.float32, dev, self.group)
                self._rows, self._vis, self._stats = rows.view(sp.cap, lay.stride), vis, stats.view(world, 4)
                self._nvls = (h_rows, h_vis, h_stats)
                self._mc_scratch = torch.zeros(sp.cap // world // 256 + 8, dtype=torch.float32, device=dev)
            else:
                self._rows = torch.zeros(sp.cap, lay.stride, dtype=torch.float32, device=dev)
                self._vis = torch.zeros(sp.cap, dtype=torch.int32, device=dev)
                self._stats = torch.zeros(world, 4, dtype=torch.float32, device=dev)
            self._buf_key = key
        return self._rows, self._vis, self._stats

    def _mark(self, name: str):
        if self.phase_events is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.phase_events.append((name, e))

    def __call__(self, viewmats: Tensor, Ks: Tensor, gt_imgs: Tensor, exposure: Optional[Tensor],
                 width: int, height: int, n_views_total: int, background: Optional[Tensor] = None,
                 step_optimizer: bool = True, gt_depths: Optional[Tensor] = None) -> MappingStepResult:
        sp = self.splats
        P = sp.params
        for p in P:
            p.grad = None
        rank, world = _world(self.group)
        if world > 1 and sp.cap % world:
            raise ValueError(f"FlatSplats(align=...) must be a multiple of the {world} ranks of the step "
                             f"(capacity {sp.cap})")
        lay = RowLayout(sp.K)
        rows, vis, stats_all = self._buffers(lay, world)
        c_local = viewmats.shape[0]
        bg = background if background is not None else torch.zeros(c_local, 3, device=viewmats.device)
        kw = {} if self.sh_degree is None else {"sh_degree": self.sh_degree}
        self._mark("begin")
        if c_local == 0:
            # more ranks than views in the window (the first keyframes of a sequence): this rank renders
            # nothing, contributes zero rows and still takes part in every collective
            from types import SimpleNamespace
            dev = sp.flat.device
            rows[:sp.n].zero_()
            zero = torch.zeros((), device=dev)
            total_share, photo, share = zero, zero, 0.0
            if gt_depths is not None:           # the window's valid-pixel count is a collective: take part
                allreduce_sum_(torch.zeros(1, device=dev), self.group)
            out = SimpleNamespace(radii=torch.zeros(0, sp.n, dtype=torch.int32, device=dev),
                                  n_touched=torch.zeros(0, sp.n, dtype=torch.int64, device=dev),
                                  depthmaps=torch.zeros(0, height, width, device=dev), means2d=None,
                                  width=width, height=height, n_cameras=0)
        else:
            out = self.render_fn(P[0], P[1], P[2], P[3], P[4], viewmats, Ks, width, height, packed=False,
                                 render_mode="RGB+D", log_uncertainties=P[5], backgrounds=bg, **kw)
            # the image losses are means over ALL views of the window: weight this rank's share
            share = c_local / float(n_views_total)
            if self._fused_loss:
                # the share is folded into the loss kernel's coefficients: `loss` is already this rank's part and
                # its gradient needs no scaling pass (photo stays the plain photometric value)
                loss, photo = self.loss_fn(out.render, gt_imgs, exposure, self.ssim_weight, "mapping", scale=share)
                total_share = loss.detach()
            else:
                total, photo = self.loss_fn(out.render, gt_imgs, exposure, self.ssim_weight, "mapping")
                loss = total * share
                total_share = total.detach() * share
            if self.tv_w:
                # a sum (not a mean) over this rank's views: no share factor (gslam/backend.py:297-318)
                loss = loss + self.tv_w * edge_aware_tv(out.depthmaps, out.rgbs, out.alphas[..., 0] > 0.4)
            if gt_depths is not None:
                # use_gt_depths (gslam/backend.py:320-323): 0.1 * mean |depth - gt| over gt > 0, a mean over
                # the window's valid pixels: this rank's share is its valid-pixel fraction of the window total
                valid = (gt_depths > 0).to(out.depthmaps.dtype)
                n_valid = allreduce_sum_(valid.sum().view(1).clone(), self.group).clamp_min(1.0)
                loss = loss + 0.1 * ((out.depthmaps - gt_depths).abs() * valid).sum() / n_valid[0]
            if self.retain_means2d_grad:
                out.means2d.retain_grad()
            kernels_write_rows = self.collect_grads is None
            if kernels_write_rows:
                ops.grad_rows_out, ops.grad_rows_written = rows, False
            try:
                loss.backward()
            finally:
                ops.grad_rows_out = None
            if kernels_write_rows:
                if not ops.grad_rows_written:
                    raise RuntimeError("MappingStep: the backward pass did not write the gradient rows "
                                       "(render_fn is not the CUDA rasterizer? pass collect_grads)")
            else:
                self.collect_grads(P, out.radii, rows, lay)
        self._mark("backward_done")

        n = sp.n
        per = sp.cap // world
        n0 = min(rank * per, n)
        n1 = min(n0 + per, n) if rank * per < n else n0
        mine = torch.stack([total_share, photo.detach() * share])
        if not step_optimizer:
            # densification decides from the whole reduced gradient on every rank and the parameters it
            # re-creates take no step: all-reduce, no update (apply_pending() runs it if nothing changed)
            allreduce_sum_(rows, self.group)
            allreduce_sum_(mine, self.group)
            vis_count = rows[:n, lay.vis].to(torch.int32)
            self._pending = (rows, lay, vis)
            return MappingStepResult(mine[0], mine[1], out.radii, out.n_touched, vis_count, out)
        self._pending = None
        if world > 1 and self._nvls is not None:
            # ONE kernel = reduce-scatter (in-switch sums of the rows) + update + all-gather (multicast stores),
            # between two cross-GPU barriers: every rank's rows written before, every store visible after
            h_rows, h_vis, h_stats = self._nvls
            h_rows.barrier()
            self._mark("reduce_scatter_done")
            self.opt.step_rows_mc(P, sp, rows, h_rows.multicast_ptr, lay, rank * per, min((rank + 1) * per, n)
                                  if rank * per < n else rank * per, h_vis.multicast_ptr, mine, h_stats.multicast_ptr,
                                  rank, self._mc_scratch, scales_index=2, iso_weight=self.iso_w, decay_index=3,
                                  decay=self.decay)
            self._mark("update_done")
            h_rows.barrier()
            tot = stats_all.sum(dim=0)
            self._mark("all_gather_done")
            return MappingStepResult(tot[0] + self.iso_w * tot[2], tot[1], out.radii, out.n_touched, vis[:n], out)
        if world > 1:
            chunk = rows[rank * per:(rank + 1) * per]
            if dist.get_backend(self.group) == "nccl":
                dist.reduce_scatter_tensor(chunk, rows, group=self.group)        # in place
            else:
                tmp = torch.empty_like(chunk)
                dist.reduce_scatter_tensor(tmp, rows.clone(), group=self.group)
                chunk.copy_(tmp)
        self._mark("reduce_scatter_done")
        iso = self.update_fn(self, rows, lay, n0, n1, vis)
        self._mark("update_done")
        if world > 1:
            nccl = dist.get_backend(self.group) == "nccl"
            src = (lambda t: t) if nccl else (lambda t: t.clone())
            entry = torch.cat([mine, iso.view(1).to(mine.dtype), mine.new_zeros(1)])
            with _coalesced(self.group, sp.flat.device):
                for blk, w in zip(sp.blocks(), sp.widths):
                    dist.all_gather_into_tensor(blk, src(blk[rank * per * w:(rank + 1) * per * w]), group=self.group)
                dist.all_gather_into_tensor(vis, src(vis[rank * per:(rank + 1) * per]), group=self.group)
                dist.all_gather_into_tensor(stats_all.view(-1), entry, group=self.group)
            tot = stats_all.sum(dim=0)          # fixed order: identical on every rank
            total_all, photo_all, iso_all = tot[0], tot[1], tot[2]
        else:
            total_all, photo_all, iso_all = mine[0], mine[1], iso.view(1)[0]
        self._mark("all_gather_done")
        return MappingStepResult(total_all + self.iso_w * iso_all, photo_all, out.radii, out.n_touched,
                                 vis[:n], out)

    def apply_pending(self) -> None:
        """After a `step_optimizer=False` call whose densification changed nothing: the update of that
        step, from the all-reduced rows, over the whole range on every rank (identical everywhere)."""
        if self._pending is None:
            raise RuntimeError("MappingStep.apply_pending(): no step_optimizer=False call is pending")
        rows, lay, vis = self._pending
        self._pending = None
        self.update_fn(self, rows, lay, 0, self.splats.n, vis)
