"""Drop-in for reference ``gslam/rasterization.py`` backed by libgsb.so.

Same call surface (``rasterization(...) -> RasterizationOutput``,
``get_new_splat_depth``; reference gslam/rasterization.py:17-71, :363-383) for
the argument subset its callers use (gslam/map.py:59-74, :88-103,
gslam/insertion.py:252); anything else raises NotImplementedError.  There is
no CPU path and no fallback.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import ClassVar, Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from . import ops


@dataclass
class RasterizationOutput:
    """Field order matters: gslam/backend.py:619 builds it positionally."""
    per_gaussian_params: ClassVar[tuple[str]] = ('radii', 'means2d')
    rgbs: torch.Tensor = None
    alphas: torch.Tensor = None
    depthmaps: torch.Tensor = None
    betas: torch.Tensor = None
    tile_width: int = None
    tile_height: int = None
    tiles_per_gauss: torch.Tensor = None
    isect_ids: torch.Tensor = None
    flatten_ids: torch.Tensor = None
    isect_offsets: torch.Tensor = None
    width: int = None
    height: int = None
    tile_size: int = None
    n_cameras: int = None
    camera_ids: torch.Tensor = None
    gaussian_ids: torch.Tensor = None
    radii: torch.Tensor = None
    means2d: torch.Tensor = None
    depths: torch.Tensor = None
    conics: torch.Tensor = None
    opacities: torch.Tensor = None
    n_touched: torch.Tensor = None


_BG_TAILS: dict = {}


def _bg_tail(device, with_beta: bool) -> Tensor:
    """(depth, beta) background: (0, e) when uncertainties are rendered (reference :253), else (0, 0)."""
    key = (str(device), with_beta)
    t = _BG_TAILS.get(key)
    if t is None:
        t = torch.tensor([[0.0, math.e if with_beta else 0.0]], dtype=torch.float32, device=device)
        _BG_TAILS[key] = t
    return t


def _unsupported(what: str):
    raise NotImplementedError(
        f"gslam_b200.rasterization: {what} is outside the subset gslam uses "
        "(SURVEY.md 8b); there is no fallback path")


def rasterization(
    means: Tensor,  # [N, 3]
    quats: Tensor,  # [N, 4]
    log_scales: Tensor,  # [N, 3]
    logit_opacities: Tensor,  # [N]
    logit_colors: Tensor,  # [N, 3]
    viewmats: Tensor,  # [C, 4, 4]
    Ks: Tensor,  # [C, 3, 3]
    width: int,
    height: int,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    radius_clip: float = 0.0,
    eps2d: float = 0.3,
    packed: bool = True,
    tile_size: int = 16,
    backgrounds: Optional[Tensor] = None,
    render_mode: Literal["RGB", "D", "ED", "RGB+D", "RGB+ED"] = "RGB",
    sparse_grad: bool = False,
    absgrad: bool = False,
    rasterize_mode: Literal["classic", "antialiased"] = "classic",
    channel_chunk: int = 32,
    camera_model: Literal["pinhole", "ortho", "fisheye"] = "pinhole",
    covars: Optional[Tensor] = None,
    log_uncertainties: Optional[Tensor] = None,
    visibility_min_T: float = 0.5,
    mask: Optional[Tensor] = None,
    sh_degree: Optional[int] = None,
    deterministic: Optional[bool] = None,
) -> RasterizationOutput:
    """Rasterize N Gaussians to C image planes; see the reference docstring
    (gslam/rasterization.py:72-119) for the argument meanings.

    ``sh_degree`` is the one extension beyond the reference surface (SURVEY 8a "SH note", BASELINE
    cfg4): None keeps the reference's ``sigmoid(logit_colors)``; 0..3 reads ``logit_colors`` as real
    spherical-harmonics coefficients ``[N, K >= (sh_degree+1)^2, 3]`` and colours each (view,
    Gaussian) pair ``max(SH(normalize(mean - camera centre)) + 0.5, 0)``, gradients to the
    coefficients and to the means.

    ``deterministic`` (None: ``ops.deterministic``): the backward pass accumulates in a fixed order
    instead of with float atomics, so gradients are bit-identical from run to run (include/gsb.h)."""
    N = means.shape[0]
    C = viewmats.shape[0]
    assert means.shape == (N, 3), means.shape
    assert quats.shape == (N, 4), quats.shape
    assert log_scales.shape == (N, 3), log_scales.shape
    assert logit_opacities.shape == (N,), logit_opacities.shape
    assert viewmats.shape == (C, 4, 4), viewmats.shape
    assert Ks.shape == (C, 3, 3), Ks.shape
    assert render_mode in ["RGB", "D", "ED", "RGB+D", "RGB+ED"], render_mode
    if packed:
        _unsupported("packed=True (use get_new_splat_depth for the packed projection)")
    if covars is not None:
        _unsupported("covars")
    if tile_size != ops.TILE:
        _unsupported(f"tile_size={tile_size}")
    if render_mode in ("ED", "RGB+ED"):
        _unsupported(f"render_mode={render_mode!r} (raises KeyError in the reference too, :344)")
    if sparse_grad or absgrad:
        _unsupported("sparse_grad / absgrad")
    if rasterize_mode != "classic" or camera_model != "pinhole":
        _unsupported("rasterize_mode / camera_model other than classic / pinhole")
    if sh_degree is None and (logit_colors.dim() != 2 or logit_colors.shape != (N, 3)):
        _unsupported(f"logit_colors of shape {tuple(logit_colors.shape)} (only [N,3] without sh_degree)")
    if not means.is_cuda:
        raise RuntimeError("gslam_b200.rasterization needs CUDA tensors; there is no CPU path")

    tile_width = math.ceil(width / float(tile_size))
    tile_height = math.ceil(height / float(tile_size))

    zeros = ops.ZeroPool(C, N, tile_width * tile_height, means.device)

    # K8+K1 (+ tile counting): activations, projection and record packing in one kernel
    radii, means2d, rec, tiles_per_gauss, tile_counts = ops.project(
        means, quats, log_scales, logit_opacities, logit_colors, log_uncertainties, viewmats, Ks,
        width, height, eps2d, near_plane, far_plane, radius_clip, tile_counts=zeros.tile_counts,
        sh_degree=sh_degree, det=deterministic)

    # K4+K5: counting sort by tile + per-tile depth sort (bit-identical to the upstream
    # emit / stable radix sort / offset-encode chain, which ops.isect_tiles still provides)
    isect_ids, flatten_ids, offsets = ops.bin_tiles(means2d, radii, rec, tile_counts, tile_width,
                                                    tile_height, overflow=zeros.overflow)
    isect_offsets = offsets[:-1].view(C, tile_height, tile_width)

    # backgrounds [C,3] -> [C,5] = (r, g, b, 0, e)   (reference :236-255)
    bg5 = None
    if backgrounds is not None:
        tail = _bg_tail(means.device, log_uncertainties is not None).expand(C, 2)
        rgb = backgrounds if render_mode != "D" else torch.zeros_like(backgrounds)
        bg5 = torch.cat([rgb, tail], dim=1)        # one kernel

    # K6
    render, alphas, n_touched = ops.composite(rec, means2d, flatten_ids, offsets, bg5, width,
                                              height, visibility_min_T, n_touched=zeros.n_touched,
                                              det=deterministic, radii=radii, isect_ids=isect_ids)

    out = RasterizationOutput(
        rgbs=render[..., :3] if render_mode not in ("D",) else None,
        alphas=alphas,
        depthmaps=render[..., 3] if render_mode in ("D", "RGB+D") else None,
        betas=render[..., 4] if log_uncertainties is not None else None,
        tile_width=tile_width,
        tile_height=tile_height,
        tiles_per_gauss=tiles_per_gauss,
        isect_ids=isect_ids,
        flatten_ids=flatten_ids,
        isect_offsets=isect_offsets,
        width=width,
        height=height,
        tile_size=tile_size,
        n_cameras=C,
        camera_ids=None,
        gaussian_ids=None,
        radii=radii,
        means2d=means2d,
        depths=rec[..., 9],
        conics=rec[..., 3:6],
        opacities=rec[..., 2],
        n_touched=n_touched,
    )
    out.render = render   # [C,H,W,5] (r g b depth beta); input of losses.photometric_ssim_loss
    return out


def get_new_splat_depth(
    new_params: dict[str, torch.Tensor],
    viewmats: Tensor,  # [C, 4, 4]
    Ks: Tensor,  # [C, 3, 3]
    width: int,
    height: int,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    radius_clip: float = 0.0,
    eps2d: float = 0.3,
    packed: bool = True,
    sparse_grad: bool = False,
    rasterize_mode: Literal["classic", "antialiased"] = "classic",
    camera_model: Literal["pinhole", "ortho", "fisheye"] = "pinhole",
    covars: Optional[Tensor] = None,
):
    """Projection only (reference gslam/rasterization.py:363-448); 'scales' are log-scales (:386-387)."""
    if covars is not None or rasterize_mode != "classic" or camera_model != "pinhole" or sparse_grad:
        _unsupported("covars / antialiased / non-pinhole / sparse_grad")
    means = new_params['means']
    quats = new_params['quats']
    log_scales = new_params['scales']
    N = means.shape[0]
    zeros = torch.zeros(N, dtype=torch.float32, device=means.device)
    with torch.no_grad():
        radii, means2d, rec, _, _ = ops.project(
            means, quats, log_scales, zeros, zeros[:, None].expand(N, 3), None, viewmats, Ks,
            width, height, eps2d, near_plane, far_plane, radius_clip)
        if packed:
            return ops.pack_visible(radii, means2d, rec)
        return radii, means2d, rec[..., 9]


def quat_scale_to_covar_preci(quats: Tensor, scales: Tensor, compute_covar: bool = True,
                              compute_preci: bool = True, triu: bool = False):
    """Shim of ``gsplat.quat_scale_to_covar_preci`` (reference gslam/insertion.py:88-91; a few
    thousand rows every 200 steps, so plain torch ops on the caller's device, not a kernel):
    covar = R diag(s^2) R^T, preci = R diag(s^-2) R^T with R from the normalised wxyz quaternion."""
    if triu:
        _unsupported("quat_scale_to_covar_preci(triu=True)")
    q = torch.nn.functional.normalize(quats, dim=-1)
    w, x, y, z = q.unbind(-1)
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=-1).reshape(q.shape[:-1] + (3, 3))
    covar = preci = None
    if compute_covar:
        M = R * scales[..., None, :]
        covar = M @ M.transpose(-1, -2)
    if compute_preci:
        P = R * (1.0 / scales[..., None, :])
        preci = P @ P.transpose(-1, -2)
    return covar, preci
