"""Autograd operators over libgsb.so.

Two autograd nodes, as the reference needs (``out.means2d.retain_grad()`` at
gslam/backend.py:326 makes means2d a real intermediate):

    project   : Gaussian params + cameras -> radii, means2d, rec, tiles_per_gauss
    composite : rec (+ means2d for gradient routing) -> render, alphas, n_touched

``rec`` is the 12-float splat record documented in include/gsb.h.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import Tensor

from . import _lib
from ._lib import call, check, lib, ptr, stream

REC = 12
CH = 5
TILE = 16
CSTRIDE = 32   # GSB_CSTRIDE: one per-tile counter per 128-byte line

# bench.py sets this to a zeroed int64[1] CUDA tensor to collect P (pairs evaluated); None otherwise
pair_counter: Optional[Tensor] = None

# Row-major gradient record of the mapping step.  When `grad_rows_out` is set (a zero-padded float32
# [N_pad >= N, row_stride] tensor), _Project.backward writes the six Gaussian gradients AND the per-Gaussian
# visibility count into it, one contiguous row per Gaussian, and hands autograd no gradient for the six
# parameters (`grad_rows_written` is raised instead): the rows of a Gaussian range are one contiguous block,
# which is what a reduce-scatter over the ranks of a view-sharded step needs, and the fused update kernel
# (gsb_map_update) reads the whole gradient of a Gaussian with four 16-byte loads.
grad_rows_out: Optional[Tensor] = None
grad_rows_written: bool = False


class RowLayout:
    """Columns of one gradient row for colours of 3 * K floats per Gaussian (K = 1: the reference's
    [N,3] logit colours; K = (deg+1)^2 .. : SH coefficients [N,K,3]).  quats first: 16-byte aligned."""

    def __init__(self, K: int = 1):
        self.K = int(K)
        self.quats, self.means, self.opacities, self.scales, self.unc, self.colors = 0, 4, 7, 8, 11, 12
        self.vis = 12 + 3 * self.K
        self.stride = -(-(self.vis + 1) // 4) * 4
        # in PARAM_NAMES order: means, quats, scales, opacities, colors, log_uncertainties
        self.cols = (self.means, self.quats, self.scales, self.opacities, self.colors, self.unc)
        self.widths = (3, 4, 3, 1, 3 * self.K, 1)


# Deterministic accumulation (bit-identical gradients and loss values from run to run): the float
# atomics of the backward pass -- v_rec in gsb_composite_bwd, v_viewmats in gsb_project_bwd, the loss
# and exposure sums of gsb_photo_loss / gsb_ssim_fwd / gsb_isotropic_reg -- are replaced by stored
# partials summed in a fixed order (include/gsb.h, "Deterministic mode").  Selected per call
# (`rasterization(..., deterministic=True)`, `photometric_ssim_loss(..., deterministic=True)`, the
# `deterministic` fields of TrackingConfig / MapConfig) or for a region with `deterministic_mode()`;
# this module attribute is the default of calls that do not say.
deterministic: bool = False


class deterministic_mode:
    """`with ops.deterministic_mode(True): ...` -- default of every call inside the block."""

    def __init__(self, flag: bool = True):
        self.flag = bool(flag)

    def __enter__(self):
        global deterministic
        self.prev, deterministic = deterministic, self.flag
        return self

    def __exit__(self, *exc):
        global deterministic
        deterministic = self.prev
        return False


def det_scratch(n_floats: int, device) -> Optional[Tensor]:
    """Uninitialised scratch for a deterministic-mode reduction (a fresh tensor: under CUDA-graph
    capture it becomes part of the graph's pool)."""
    return torch.empty(max(int(n_floats), 1), dtype=torch.float32, device=device)


def _c(t: Optional[Tensor]) -> Optional[Tensor]:
    if t is None:
        return None
    if t.dtype != torch.float32:
        raise TypeError(f"gslam_b200 kernels are float32 only, got {t.dtype}")
    return t.contiguous()


def tile_bits(tile_w: int, tile_h: int, n_cameras: int):
    return (int(math.floor(math.log2(tile_w * tile_h))) + 1,
            int(math.floor(math.log2(n_cameras))) + 1)


class _Project(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats, Ks,
                width, height, eps2d, near_plane, far_plane, radius_clip, tile_w, tile_h, tile_counts,
                sh_degree=None, det=False):
        """sh_degree None: logit_colors [N,3] -> sigmoid (the reference).  sh_degree 0..3 (extension):
        logit_colors holds SH coefficients [N,K,3], K >= (sh_degree+1)^2."""
        ctx.set_materialize_grads(False)    # no zero-fill kernels for the integer outputs' "gradients"
        means, quats, log_scales = _c(means), _c(quats), _c(log_scales)
        logit_opacities, logit_colors, log_unc = _c(logit_opacities), _c(logit_colors), _c(log_unc)
        viewmats, Ks = _c(viewmats), _c(Ks)
        N, C = means.shape[0], viewmats.shape[0]
        dev = means.device
        radii = torch.empty(C, N, dtype=torch.int32, device=dev)
        tiles = torch.empty(C, N, dtype=torch.int32, device=dev)
        means2d = torch.empty(C, N, 2, dtype=torch.float32, device=dev)
        rec = torch.empty(C, N, REC, dtype=torch.float32, device=dev)
        if tile_counts is None:
            tile_counts = torch.zeros(C, tile_h * tile_w, CSTRIDE, dtype=torch.int32, device=dev)
        sh = sh_degree is not None
        if sh:
            if not (0 <= sh_degree <= 3):
                raise NotImplementedError("sh_degree must be 0..3")
            if logit_colors.dim() != 3 or logit_colors.shape[2] != 3 or \
                    logit_colors.shape[1] < (sh_degree + 1) ** 2:
                raise ValueError("with sh_degree set, logit_colors must be [N, K >= (sh_degree+1)^2, 3]")
        call("gsb_project_fwd",
            ptr(means), ptr(quats), ptr(log_scales), ptr(logit_opacities),
            None if sh else ptr(logit_colors),
            ptr(log_unc), ptr(viewmats), ptr(Ks), C, N, width, height, eps2d, near_plane, far_plane,
            radius_clip, tile_w, tile_h, ptr(radii), ptr(means2d), ptr(rec), ptr(tiles),
            ptr(tile_counts), ptr(logit_colors) if sh else None, sh_degree if sh else -1,
            logit_colors.shape[1] if sh else 0, stream())
        ctx.sh_degree = sh_degree
        ctx.det = bool(det)
        ctx.save_for_backward(means, quats, log_scales, logit_opacities, logit_colors, log_unc,
                              viewmats, Ks, radii, rec)
        ctx.dims = (C, N, width, height, eps2d)
        ctx.mark_non_differentiable(radii, tiles, tile_counts)
        return radii, means2d, rec, tiles, tile_counts

    @staticmethod
    def backward(ctx, _g_radii, g_means2d, g_rec, _g_tiles, _g_counts):
        (means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats, Ks, radii,
         rec) = ctx.saved_tensors
        C, N, width, height, eps2d = ctx.dims
        dev = means.device
        if g_rec is None:
            g_rec = torch.zeros(C, N, REC, dtype=torch.float32, device=dev)
        g_rec = g_rec.contiguous()
        if g_means2d is None:
            g_means2d = torch.zeros(C, N, 2, dtype=torch.float32, device=dev)
        # a [..., 0:2] view of a record-shaped tensor is read in place (stride 12)
        if g_means2d.stride() == (N * REC, REC, 1):
            m2_stride = REC
        else:
            g_means2d = g_means2d.contiguous()
            m2_stride = 2
        need = ctx.needs_input_grad
        need_gauss = any(need[:6])
        need_view = need[6]
        sh = ctx.sh_degree is not None
        rows = grad_rows_out if need_gauss else None
        if rows is not None:
            # the mapping step's row layout: everything goes into the caller's [N_pad, stride] record
            global grad_rows_written
            lay = RowLayout(logit_colors.shape[1] if sh else 1)
            if not (rows.is_cuda and rows.dtype == torch.float32 and rows.is_contiguous() and rows.dim() == 2
                    and rows.shape[0] >= N and rows.shape[1] == lay.stride):
                raise ValueError(f"ops.grad_rows_out must be a contiguous float32 [>= {N}, {lay.stride}] CUDA tensor")
            v_view = torch.zeros(C, 4, 4, dtype=torch.float32, device=dev) if need_view else None
            det_part = det_scratch(lib().gsb_project_bwd_det_floats(C, N), dev) if (ctx.det and need_view) else None
            base = rows.data_ptr()
            col = lambda c: base + 4 * c   # noqa: E731
            call("gsb_project_bwd",
                 ptr(means), ptr(quats), ptr(log_scales), ptr(logit_opacities),
                 None if sh else ptr(logit_colors),
                 ptr(log_unc), ptr(viewmats), ptr(Ks), C, N, width, height, eps2d, ptr(radii), ptr(rec),
                 ptr(g_means2d), m2_stride, ptr(g_rec), 1, col(lay.means), col(lay.quats), col(lay.scales),
                 col(lay.opacities), None if sh else col(lay.colors), col(lay.unc), ptr(v_view), ptr(det_part),
                 lay.stride, col(lay.vis), stream())
            if sh:
                call("gsb_sh_bwd", ptr(means), ptr(logit_colors), ctx.sh_degree, logit_colors.shape[1],
                     ptr(viewmats), ptr(radii), ptr(g_rec), C, N, col(lay.colors), col(lay.means), lay.stride,
                     stream())
            grad_rows_written = True
            return (None, None, None, None, None, None, v_view, None, None, None, None, None, None, None, None,
                    None, None, None, None)
        if need_gauss:
            # one flat buffer, means|quats|scales|opacities|colors|log_unc, so that a
            # view-sharded mapping step can all-reduce the six gradients in ONE call
            flat = torch.empty(N * 15, dtype=torch.float32, device=dev)
            v_means = flat[0:3 * N].view(N, 3)
            v_quats = flat[3 * N:7 * N].view(N, 4)
            v_scales = flat[7 * N:10 * N].view(N, 3)
            v_opac = flat[10 * N:11 * N]
            v_colors = flat[11 * N:14 * N].view(N, 3)
            v_unc = flat[14 * N:15 * N]
            if log_unc is None:
                v_unc.zero_()
        else:
            v_means = v_quats = v_scales = v_opac = v_colors = v_unc = None
        v_view = torch.zeros(C, 4, 4, dtype=torch.float32, device=dev) if need_view else None
        det_part = None
        if ctx.det and need_view:
            det_part = det_scratch(lib().gsb_project_bwd_det_floats(C, N), dev)
        if need_gauss or need_view:
            call("gsb_project_bwd",
                ptr(means), ptr(quats), ptr(log_scales), ptr(logit_opacities),
                None if sh else ptr(logit_colors),
                ptr(log_unc), ptr(viewmats), ptr(Ks), C, N, width, height, eps2d, ptr(radii), ptr(rec),
                ptr(g_means2d), m2_stride, ptr(g_rec), int(need_gauss), ptr(v_means), ptr(v_quats),
                ptr(v_scales), ptr(v_opac), None if sh else ptr(v_colors), ptr(v_unc), ptr(v_view),
                ptr(det_part), 0, None, stream())
        if sh and need_gauss:
            v_colors = torch.empty_like(logit_colors)
            call("gsb_sh_bwd", ptr(means), ptr(logit_colors), ctx.sh_degree, logit_colors.shape[1],
                 ptr(viewmats), ptr(radii), ptr(g_rec), C, N, ptr(v_colors), ptr(v_means), 0, stream())
        grads = [v_means, v_quats, v_scales, v_opac, v_colors,
                 v_unc if log_unc is not None else None]
        grads = [g if n else None for g, n in zip(grads, need[:6])]
        return (*grads, v_view, None, None, None, None, None, None, None, None, None, None, None, None)


class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rec, means2d, flatten_ids, offsets, backgrounds, width, height, tile_w, tile_h,
                visibility_min_T, count_touched, n_touched, det=False, radii=None, isect_ids=None):
        ctx.set_materialize_grads(False)
        # means2d is not read (the record carries x,y); it is an input only so that its
        # gradient is delivered as a separate autograd edge.
        C, N = rec.shape[0], rec.shape[1]
        dev = rec.device
        # offsets with C*T+1 entries carry the intersection count on the device (bin_tiles)
        n_isects = -(flatten_ids.numel() + 1) if offsets.numel() == C * tile_w * tile_h + 1 else flatten_ids.numel()
        backgrounds = _c(backgrounds)
        render = torch.empty(C, height, width, CH, dtype=torch.float32, device=dev)
        alphas = torch.empty(C, height, width, 1, dtype=torch.float32, device=dev)
        last_ids = torch.empty(C, height, width, dtype=torch.int32, device=dev)
        if not count_touched:
            n_touched = None
        elif n_touched is None:
            n_touched = torch.zeros(C, N, dtype=torch.int64, device=dev)
        call("gsb_composite_fwd", 
            ptr(rec), ptr(flatten_ids), ptr(offsets), ptr(backgrounds), C, N, width, height, tile_w,
            tile_h, n_isects, visibility_min_T, ptr(render), ptr(alphas), ptr(last_ids),
            ptr(n_touched), ptr(pair_counter), stream())
        if det and (radii is None or isect_ids is None):
            raise ValueError("deterministic compositing backward needs radii and isect_ids")
        ctx.det = bool(det)
        ctx.save_for_backward(rec, flatten_ids, offsets, backgrounds, alphas, last_ids,
                              radii if det else None, isect_ids if det else None)
        ctx.dims = (C, N, width, height, tile_w, tile_h, n_isects)
        if n_touched is None:
            n_touched = torch.zeros(C, N, dtype=torch.int64, device=dev)
        ctx.mark_non_differentiable(n_touched)
        return render, alphas, n_touched

    @staticmethod
    def backward(ctx, v_render, v_alphas, _g_touched):
        rec, flatten_ids, offsets, backgrounds, alphas, last_ids, radii, isect_ids = ctx.saved_tensors
        C, N, width, height, tile_w, tile_h, n_isects = ctx.dims
        dev = rec.device
        if v_render is None:
            v_render = torch.zeros(C, height, width, CH, dtype=torch.float32, device=dev)
        v_render = v_render.contiguous()
        v_alphas = v_alphas.contiguous() if v_alphas is not None else None
        need_bg = backgrounds is not None and ctx.needs_input_grad[4]
        cap = flatten_ids.numel()
        if ctx.det and cap > 0:
            # rows + fixed-order gather: v_rec is written for every (c, n), nothing to zero
            v_rec = torch.empty(C, N, REC, dtype=torch.float32, device=dev)
            nbytes = lib().gsb_composite_bwd_det_bytes(cap)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            v_bg = None
            call("gsb_composite_bwd",
                 ptr(rec), ptr(flatten_ids), ptr(offsets), ptr(backgrounds), C, N, width, height, tile_w,
                 tile_h, n_isects, ptr(alphas), ptr(last_ids), ptr(v_render), ptr(v_alphas),
                 ptr(v_rec), None, ptr(radii), ptr(isect_ids), ptr(ws), nbytes, stream())
            if need_bg:     # sum_pixels (1 - alpha) v_render: a plain (deterministic) torch reduction
                v_bg = ((1.0 - alphas) * v_render).sum(dim=(1, 2))
        else:
            v_rec = torch.zeros(C, N, REC, dtype=torch.float32, device=dev)
            v_bg = torch.zeros(C, CH, dtype=torch.float32, device=dev) if need_bg else None
            call("gsb_composite_bwd",
                 ptr(rec), ptr(flatten_ids), ptr(offsets), ptr(backgrounds), C, N, width, height, tile_w,
                 tile_h, n_isects, ptr(alphas), ptr(last_ids), ptr(v_render), ptr(v_alphas),
                 ptr(v_rec), ptr(v_bg), None, None, None, 0, stream())
        # x,y gradients travel on the means2d edge; _Project.backward ignores those two slots of v_rec
        return (v_rec, v_rec[..., 0:2], None, None, v_bg, None, None, None, None, None, None, None,
                None, None, None)


def project(means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats, Ks, width,
            height, eps2d=0.3, near_plane=0.01, far_plane=1e10, radius_clip=0.0, tile_counts=None,
            sh_degree=None, det=None):
    """tile_counts: optional pre-zeroed int32 [C, T, 32] (see ZeroPool); det: deterministic backward
    (None: the module default `deterministic`)."""
    tile_w, tile_h = math.ceil(width / TILE), math.ceil(height / TILE)
    return _Project.apply(means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats,
                          Ks, int(width), int(height), float(eps2d), float(near_plane),
                          float(far_plane), float(radius_clip), tile_w, tile_h, tile_counts, sh_degree,
                          deterministic if det is None else bool(det))


class ZeroPool:
    """Every zero-initialised integer buffer of one rasterization() call, cleared by ONE fill kernel:
    tile_counts int32 [C,T,32] | overflow int32 [1] | n_touched int64 [C,N]."""

    def __init__(self, C: int, N: int, T: int, device):
        n_counts = C * T * CSTRIDE
        head = n_counts + 2                      # + overflow, + pad to an 8-byte boundary
        pool = torch.zeros(head + 2 * C * N, dtype=torch.int32, device=device)
        self.tile_counts = pool[:n_counts].view(C, T, CSTRIDE)
        self.overflow = pool[n_counts:n_counts + 1]
        self.n_touched = pool[head:].view(torch.int64).view(C, N)


@torch.no_grad()
def inclusive_scan(x_i32: Tensor, flag_positive: bool = False) -> Tensor:
    n = x_i32.numel()
    out = torch.empty(n, dtype=torch.int64, device=x_i32.device)
    if n:
        nbytes = lib().gsb_scan_workspace_bytes(n)
        ws = _lib.workspace("scan", nbytes, x_i32.device)
        call("gsb_scan_i32_i64", ptr(x_i32), n, int(flag_positive), ptr(out), ptr(ws), ws.numel(),
                                     stream())
    return out


def read_i64(t: Tensor, index: int = -1) -> int:
    """The one explicit device->host sync of the path."""
    import ctypes as ct
    if t.numel() == 0:
        return 0
    host = ct.c_int64(0)
    src = t.data_ptr() + (index % t.numel()) * 8
    call("gsb_read_i64_host", src, ct.addressof(host), stream())
    return int(host.value)


@torch.no_grad()
def sort_pairs(keys: Tensor, vals: Tensor, end_bit: int):
    """Stable radix sort of (int64 key, int32 value); clobbers the inputs."""
    n = keys.numel()
    keys_out, vals_out = torch.empty_like(keys), torch.empty_like(vals)
    if n:
        nbytes = lib().gsb_sort_workspace_bytes(n)
        ws = _lib.workspace("sort", nbytes, keys.device)
        call("gsb_sort_pairs", ptr(keys), ptr(vals), ptr(keys_out), ptr(vals_out), n, end_bit,
                                   ptr(ws), ws.numel(), stream())
    return keys_out, vals_out


@torch.no_grad()
def isect_tiles(means2d, radii, rec, tiles_per_gauss, tile_w, tile_h, sort=True):
    """K4 (reference gslam/rasterization.py:261-272): -> isect_ids[I] i64, flatten_ids[I] i32."""
    C, N = radii.shape
    dev = radii.device
    tile_n_bits, cam_n_bits = tile_bits(tile_w, tile_h, C)
    cum = inclusive_scan(tiles_per_gauss.view(-1))
    n_isects = read_i64(cum)
    isect_ids = torch.empty(n_isects, dtype=torch.int64, device=dev)
    flatten_ids = torch.empty(n_isects, dtype=torch.int32, device=dev)
    if n_isects:
        call("gsb_isect_emit", ptr(means2d), ptr(radii), ptr(rec), ptr(cum), C, N, tile_w, tile_h,
                                   tile_n_bits, ptr(isect_ids), ptr(flatten_ids), stream())
        if sort:
            isect_ids, flatten_ids = sort_pairs(isect_ids, flatten_ids, 32 + tile_n_bits + cam_n_bits)
    return isect_ids, flatten_ids


@torch.no_grad()
def isect_offset_encode(isect_ids, n_cameras, tile_w, tile_h):
    """K5 (reference gslam/rasterization.py:274)."""
    tile_n_bits, _ = tile_bits(tile_w, tile_h, n_cameras)
    offsets = torch.empty(n_cameras, tile_h, tile_w, dtype=torch.int32, device=isect_ids.device)
    call("gsb_isect_offsets", ptr(isect_ids), isect_ids.numel(), n_cameras, tile_w, tile_h,
                                  tile_n_bits, ptr(offsets), stream())
    return offsets


# Capacity mode (no host sync, CUDA-graph capturable): when set, bin_tiles() sizes the key
# buffers with this many entries instead of reading the intersection count back; the
# device flag `last_overflow` is non-zero if a call produced more intersections than that.
isect_capacity: Optional[int] = None
last_overflow: Optional[Tensor] = None
last_n_isects: Optional[Tensor] = None


@torch.no_grad()
def bin_tiles(means2d, radii, rec, tile_counts, tile_w, tile_h, overflow=None):
    """K4+K5, tile-binned (csrc/binning.cu): -> isect_ids, flatten_ids, offsets[C*T+1] i32.

    Bit-identical to isect_tiles(sort=True) + isect_offset_encode.  In the default
    (strict) mode the intersection count is read back once to size the outputs exactly,
    like upstream; with `isect_capacity` set nothing synchronises."""
    global last_overflow, last_n_isects
    C, N = radii.shape
    dev = radii.device
    T = tile_w * tile_h
    tile_n_bits, _ = tile_bits(tile_w, tile_h, C)
    offsets = torch.empty(C * T + 1, dtype=torch.int32, device=dev)
    cursors = torch.empty(C * T * CSTRIDE, dtype=torch.int32, device=dev)
    n_dev = torch.empty(1, dtype=torch.int64, device=dev)
    if overflow is None:
        overflow = torch.zeros(1, dtype=torch.int32, device=dev)
    call("gsb_tile_scan", ptr(tile_counts), C, tile_w, tile_h, ptr(offsets), ptr(cursors), ptr(n_dev),
         stream())
    cap = isect_capacity if isect_capacity is not None else read_i64(n_dev)
    depth_id = torch.empty(cap, dtype=torch.int64, device=dev)
    isect_ids = torch.empty(cap, dtype=torch.int64, device=dev)
    flatten_ids = torch.empty(cap, dtype=torch.int32, device=dev)
    if cap:
        call("gsb_isect_scatter", ptr(means2d), ptr(radii), ptr(rec), C, N, tile_w, tile_h, ptr(cursors),
             cap, ptr(depth_id), ptr(overflow), stream())
        call("gsb_segment_sort", ptr(depth_id), ptr(isect_ids), ptr(offsets), C, tile_w, tile_h,
             tile_n_bits, cap,
             ptr(isect_ids), ptr(flatten_ids), stream())
    last_overflow, last_n_isects = overflow, n_dev
    return isect_ids, flatten_ids, offsets


def composite(rec, means2d, flatten_ids, offsets, backgrounds, width, height, visibility_min_T=0.5,
              count_touched=True, n_touched=None, det=None, radii=None, isect_ids=None):
    """-> render [C,H,W,5], alphas [C,H,W,1], n_touched int64 [C,N] (n_touched: optional pre-zeroed buffer).
    det (None: the module default): deterministic backward; it needs radii [C,N] and the sorted isect_ids."""
    tile_w, tile_h = math.ceil(width / TILE), math.ceil(height / TILE)
    det = deterministic if det is None else bool(det)
    return _Composite.apply(rec, means2d, flatten_ids, offsets, backgrounds, int(width), int(height),
                            tile_w, tile_h, float(visibility_min_T), bool(count_touched), n_touched,
                            det, radii if det else None, isect_ids if det else None)


@torch.no_grad()
def pack_visible(radii, means2d, rec):
    """K3 compaction (reference gslam/rasterization.py:409-419)."""
    C, N = radii.shape
    dev = radii.device
    cum = inclusive_scan(radii.view(-1), flag_positive=True)
    nnz = read_i64(cum)
    camera_ids = torch.empty(nnz, dtype=torch.int64, device=dev)
    gaussian_ids = torch.empty(nnz, dtype=torch.int64, device=dev)
    radii_out = torch.empty(nnz, dtype=torch.int32, device=dev)
    means2d_out = torch.empty(nnz, 2, dtype=torch.float32, device=dev)
    depths_out = torch.empty(nnz, dtype=torch.float32, device=dev)
    if nnz:
        call("gsb_pack_visible", ptr(radii), ptr(means2d), ptr(rec), ptr(cum), C, N,
                                     ptr(camera_ids), ptr(gaussian_ids), ptr(radii_out),
                                     ptr(means2d_out), ptr(depths_out), stream())
    return camera_ids, gaussian_ids, radii_out, means2d_out, depths_out
