"""CPU oracle for the gslam rasterization hot path (TEST INFRASTRUCTURE ONLY).

This file is a pure-PyTorch, CPU-runnable restatement of the arithmetic behind
``/root/reference/gslam/rasterization.py``.  It is the checker for the CUDA
path in ``gslam_b200/``; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.
The product path never does (it fails loudly when the CUDA library is absent).

PARITY UNPINNED.  The reference's arithmetic lives in two un-vendored
third-party CUDA extensions that are absent from /root/reference and from this
image:

* ``gsplat`` -- fork ``abhigyan7/gsplat`` branch ``gslam`` (no pinned commit,
  reference README.md:11); lineage nerfstudio-project/gsplat around commit
  795161945b37747709d4da965b226a19fdf87d3f (cited at gslam/primitives.py:39,
  gslam/insertion.py:37, gslam/pruning.py:22).
* ``fused_ssim`` -- rahul-goel/fused-ssim@30fb258c8a38fe61e640c382f891f14b2e8b0b5a
  (reference README.md:15).

The reference holds no golden vectors, fixtures or tests for this path
(SURVEY.md section 4), so the oracle restates the published gsplat algorithm
(EWA splatting, tile binning with 64-bit keys, front-to-back compositing) and
anchors on the reference's own call sites, each cited below as file:line into
/root/reference.  Its own known-answer tests live in tests/test_oracle.py.

Everything works in float32 by default and in float64 when the inputs are
float64 (used for finite-difference gradient checks).  Gradients come from
torch autograd.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import ClassVar, Optional

import torch
import torch.nn.functional as F
from torch import Tensor

ALPHA_MAX = 0.999          # gsplat rasterize_to_pixels: alpha = min(0.999, ...)
ALPHA_MIN = 1.0 / 255.0    # pairs below this are skipped
T_STOP = 1e-4              # stop when T * (1 - alpha) <= 1e-4
FOV_MARGIN = 0.3           # 0.3 * tan(fov/2) clamp of x/z, y/z in the EWA Jacobian
RADIUS_GUARD = 0.01        # max(0.01, b^2 - det) under the eigenvalue sqrt


# ----------------------------------------------------------------------------
# K8  activations  (reference gslam/rasterization.py:145-149)
# ----------------------------------------------------------------------------
def activations(log_scales, logit_opacities, logit_colors, log_uncertainties=None):
    opacities = torch.sigmoid(logit_opacities)           # :145
    colors = torch.sigmoid(logit_colors)                 # :146
    scales = torch.exp(log_scales)                       # :147
    betas = None
    if log_uncertainties is not None:
        betas = torch.exp(log_uncertainties).clamp(min=0.01)   # :149
    return scales, opacities, colors, betas


# ----------------------------------------------------------------------------
# Spherical-harmonics colours: EXTENSION (SURVEY 8a "SH note"); the reference has none.
# Published algorithm restated: real SH basis to degree 3 in the 3DGS / gsplat sign convention,
# colour = clamp_min(sum_k Y_k(d) c_k + 0.5, 0), d = normalize(mean - camera centre), the camera
# centre taken as a constant.  Written as explicit polynomials, independent of the kernel's code.
# ----------------------------------------------------------------------------
def sh_basis(degree: int, dirs: Tensor) -> Tensor:
    """dirs [..., 3] unit vectors -> [..., (degree+1)^2]."""
    x, y, z = dirs.unbind(-1)
    Y = [torch.full_like(x, 0.28209479177387814)]
    if degree >= 1:
        c1 = 0.4886025119029199
        Y += [-c1 * y, c1 * z, -c1 * x]
    if degree >= 2:
        Y += [1.0925484305920792 * x * y, -1.0925484305920792 * y * z,
              0.31539156525252005 * (2 * z * z - x * x - y * y),
              -1.0925484305920792 * x * z, 0.5462742152960396 * (x * x - y * y)]
    if degree >= 3:
        Y += [-0.5900435899266435 * y * (3 * x * x - y * y), 2.890611442640554 * x * y * z,
              -0.4570457994644658 * y * (4 * z * z - x * x - y * y),
              0.3731763325901154 * z * (2 * z * z - 3 * x * x - 3 * y * y),
              -0.4570457994644658 * x * (4 * z * z - x * x - y * y),
              1.445305721320277 * z * (x * x - y * y), -0.5900435899266435 * x * (x * x - 3 * y * y)]
    return torch.stack(Y, dim=-1)


def sh_colors(degree: int, coeffs: Tensor, means: Tensor, viewmats: Tensor) -> Tensor:
    """coeffs [N,K,3], means [N,3], viewmats [C,4,4] -> colours [C,N,3]."""
    K = (degree + 1) ** 2
    Rm, t = viewmats[:, :3, :3].detach(), viewmats[:, :3, 3].detach()
    centre = -(Rm.transpose(1, 2) @ t[..., None])[..., 0]                  # [C,3]
    dirs = F.normalize(means[None] - centre[:, None], dim=-1)               # [C,N,3]
    Y = sh_basis(degree, dirs)                                              # [C,N,K]
    return ((Y[..., None] * coeffs[None, :, :K, :]).sum(-2) + 0.5).clamp_min(0.0)


# ----------------------------------------------------------------------------
# quaternion (wxyz, not necessarily unit) + scale -> covariance
# (gsplat.quat_scale_to_covar_preci; reference call gslam/insertion.py:88)
# ----------------------------------------------------------------------------
def quat_to_rotmat(quats: Tensor) -> Tensor:
    q = F.normalize(quats, dim=-1)
    w, x, y, z = q.unbind(-1)
    R = torch.stack(
        [
            1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y),
        ],
        dim=-1,
    )
    return R.reshape(quats.shape[:-1] + (3, 3))


def quat_scale_to_covar(quats: Tensor, scales: Tensor) -> Tensor:
    R = quat_to_rotmat(quats)
    M = R * scales[..., None, :]
    return M @ M.transpose(-1, -2)


# ----------------------------------------------------------------------------
# K1  fully_fused_projection, unpacked (reference gslam/rasterization.py:153-170)
# ----------------------------------------------------------------------------
def fully_fused_projection(
    means: Tensor,      # [N,3]
    quats: Tensor,      # [N,4]
    scales: Tensor,     # [N,3] post-exp
    viewmats: Tensor,   # [C,4,4] world->camera
    Ks: Tensor,         # [C,3,3]
    width: int,
    height: int,
    eps2d: float = 0.3,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    radius_clip: float = 0.0,
):
    """Returns radii[C,N] int32, means2d[C,N,2], depths[C,N], conics[C,N,3].

    Entries with radii == 0 are culled; their float outputs are zero.
    """
    R = viewmats[:, :3, :3]                                    # [C,3,3]
    t = viewmats[:, :3, 3]                                     # [C,3]
    p = torch.einsum("cij,nj->cni", R, means) + t[:, None, :]  # [C,N,3]
    x, y, z = p.unbind(-1)
    valid = (z >= near_plane) & (z <= far_plane)
    zs = torch.where(valid, z, torch.ones_like(z))             # keep 1/z finite

    covar = quat_scale_to_covar(quats, scales)                 # [N,3,3]
    covar_c = torch.einsum("cij,njk,clk->cnil", R, covar, R)   # [C,N,3,3]

    fx, fy = Ks[:, 0, 0][:, None], Ks[:, 1, 1][:, None]
    cx, cy = Ks[:, 0, 2][:, None], Ks[:, 1, 2][:, None]
    tan_fovx = 0.5 * width / fx
    tan_fovy = 0.5 * height / fy
    lim_xp = (width - cx) / fx + FOV_MARGIN * tan_fovx
    lim_xn = cx / fx + FOV_MARGIN * tan_fovx
    lim_yp = (height - cy) / fy + FOV_MARGIN * tan_fovy
    lim_yn = cy / fy + FOV_MARGIN * tan_fovy
    rz = 1.0 / zs
    tx = zs * torch.minimum(lim_xp, torch.maximum(-lim_xn, x * rz))
    ty = zs * torch.minimum(lim_yp, torch.maximum(-lim_yn, y * rz))
    zero = torch.zeros_like(rz)
    J = torch.stack(
        [fx * rz, zero, -fx * tx * rz * rz, zero, fy * rz, -fy * ty * rz * rz], dim=-1
    ).reshape(p.shape[:2] + (2, 3))
    cov2d = J @ covar_c @ J.transpose(-1, -2)                  # [C,N,2,2]
    mx = fx * x * rz + cx
    my = fy * y * rz + cy

    c00 = cov2d[..., 0, 0] + eps2d
    c01 = 0.5 * (cov2d[..., 0, 1] + cov2d[..., 1, 0])
    c11 = cov2d[..., 1, 1] + eps2d
    det = c00 * c11 - c01 * c01
    valid = valid & (det > 0)
    dets = torch.where(valid, det, torch.ones_like(det))
    conics = torch.stack([c11 / dets, -c01 / dets, c00 / dets], dim=-1)

    b = 0.5 * (c00 + c11)
    v1 = b + torch.sqrt(torch.clamp(b * b - det, min=RADIUS_GUARD))
    radius = torch.ceil(3.0 * torch.sqrt(v1))
    valid = valid & (radius > radius_clip)
    valid = valid & ~(
        (mx + radius <= 0) | (mx - radius >= width) | (my + radius <= 0) | (my - radius >= height)
    )
    radii = torch.where(valid, radius, torch.zeros_like(radius)).to(torch.int32)
    vf = valid.to(means.dtype)
    means2d = torch.stack([mx, my], dim=-1) * vf[..., None]
    depths = z * vf
    conics = conics * vf[..., None]
    return radii, means2d, depths, conics


# ----------------------------------------------------------------------------
# K4  isect_tiles  (reference gslam/rasterization.py:261-272)
# ----------------------------------------------------------------------------
def tile_bits(tile_width: int, tile_height: int, n_cameras: int):
    """Bit widths of the tile and camera fields of an intersection key."""
    n_tiles = tile_width * tile_height
    tile_n_bits = int(math.floor(math.log2(n_tiles))) + 1
    cam_n_bits = int(math.floor(math.log2(n_cameras))) + 1
    return tile_n_bits, cam_n_bits


def tile_bbox(means2d: Tensor, radii: Tensor, tile_size: int, tile_width: int, tile_height: int):
    """[floor((x-r)/ts), ceil((x+r)/ts)) clamped to the tile grid, float32 as on the device."""
    m = means2d.to(torch.float32)
    tr = radii.to(torch.float32) / tile_size
    tx = m[..., 0] / tile_size
    ty = m[..., 1] / tile_size
    x0 = torch.floor(tx - tr).clamp(0, tile_width).to(torch.int64)
    x1 = torch.ceil(tx + tr).clamp(0, tile_width).to(torch.int64)
    y0 = torch.floor(ty - tr).clamp(0, tile_height).to(torch.int64)
    y1 = torch.ceil(ty + tr).clamp(0, tile_height).to(torch.int64)
    return x0, x1, y0, y1


@torch.no_grad()
def isect_tiles(
    means2d: Tensor,   # [C,N,2]
    radii: Tensor,     # [C,N] int32
    depths: Tensor,    # [C,N]
    tile_size: int,
    tile_width: int,
    tile_height: int,
    sort: bool = True,
):
    """Returns tiles_per_gauss[C,N] i32, isect_ids[I] i64 (sorted), flatten_ids[I] i32.

    key = cam << (32 + tile_n_bits) | tile << 32 | bits(float32 depth); value =
    c * N + n.  Emission order is c-major, then n, then tile rows, then tile
    columns; the sort is stable so equal keys keep that order.
    """
    C, N = radii.shape
    tile_n_bits, _ = tile_bits(tile_width, tile_height, C)
    x0, x1, y0, y1 = tile_bbox(means2d, radii, tile_size, tile_width, tile_height)
    live = radii > 0
    cnt = torch.where(live, (x1 - x0) * (y1 - y0), torch.zeros_like(x0))
    tiles_per_gauss = cnt.to(torch.int32)

    flat_cnt = cnt.reshape(-1)
    total = int(flat_cnt.sum())
    gid = torch.repeat_interleave(torch.arange(C * N), flat_cnt)           # [I]
    starts = torch.cumsum(flat_cnt, 0) - flat_cnt
    local = torch.arange(total) - starts[gid]
    bw = (x1 - x0).reshape(-1)[gid]
    tyy = y0.reshape(-1)[gid] + local // torch.clamp(bw, min=1)
    txx = x0.reshape(-1)[gid] + local % torch.clamp(bw, min=1)
    tile_id = tyy * tile_width + txx
    cam = gid // N
    dbits = depths.to(torch.float32).contiguous().view(torch.int32).reshape(-1)[gid].to(torch.int64)
    dbits = dbits & 0xFFFFFFFF
    isect_ids = (cam << (32 + tile_n_bits)) | (tile_id << 32) | dbits
    flatten_ids = gid.to(torch.int32)
    if sort:
        isect_ids, order = torch.sort(isect_ids, stable=True)
        flatten_ids = flatten_ids[order]
    return tiles_per_gauss, isect_ids, flatten_ids


# ----------------------------------------------------------------------------
# K5  isect_offset_encode  (reference gslam/rasterization.py:274)
# ----------------------------------------------------------------------------
@torch.no_grad()
def isect_offset_encode(isect_ids: Tensor, n_cameras: int, tile_width: int, tile_height: int):
    """offsets[c,ty,tx] = index of the first sorted key of (c, tile); an empty
    tile gets the start of the next non-empty one (n_isects after the last)."""
    n_tiles = tile_width * tile_height
    tile_n_bits, _ = tile_bits(tile_width, tile_height, n_cameras)
    cam = isect_ids >> (32 + tile_n_bits)
    tile = (isect_ids >> 32) & ((1 << tile_n_bits) - 1)
    lin = cam * n_tiles + tile
    q = torch.arange(n_cameras * n_tiles, dtype=torch.int64)
    off = torch.searchsorted(lin.contiguous(), q, right=False)
    return off.to(torch.int32).reshape(n_cameras, tile_height, tile_width)


# ----------------------------------------------------------------------------
# K6/K7  rasterize_to_pixels (fork variant, reference gslam/rasterization.py:325-339)
# ----------------------------------------------------------------------------
def rasterize_to_pixels(
    means2d: Tensor,        # [C,N,2]
    conics: Tensor,         # [C,N,3]
    colors: Tensor,         # [C,N,D]
    opacities: Tensor,      # [C,N]
    width: int,
    height: int,
    tile_size: int,
    isect_offsets: Tensor,  # [C,th,tw] int32
    flatten_ids: Tensor,    # [I] int32
    backgrounds: Optional[Tensor] = None,   # [C,D]
    visibility_min_T: float = 0.5,
    return_stats: bool = False,
):
    """Front-to-back alpha compositing, one 16x16 tile at a time.

    The sequential per-pixel loop of the CUDA reference
        sigma = .5(a dx^2 + c dy^2) + b dx dy;  alpha = min(.999, o e^-sigma)
        skip if sigma < 0 or alpha < 1/255;  stop if T(1-alpha) <= 1e-4
        out += color * alpha * T;  T *= 1 - alpha
    is vectorised per tile with an inclusive cumulative product: a pair is kept
    iff it passes the skip test and the running product including it is still
    > 1e-4 (the product is monotone, so this equals the sequential break).

    n_touched[c,n] (fork extra; definition adopted in SURVEY.md 8a/a7) counts
    the pixels where Gaussian n was accumulated while the transmittance in
    front of it was > visibility_min_T.

    Returns render_colors[C,H,W,D], render_alphas[C,H,W,1], n_touched[C,N] int32
    (+ dict(last_ids[C,H,W] int32, P=int) when return_stats).
    """
    C, N = opacities.shape
    D = colors.shape[-1]
    dt = means2d.dtype
    th, tw = isect_offsets.shape[1:]
    n_isects = flatten_ids.numel()
    off = isect_offsets.reshape(-1).tolist() + [n_isects]
    fid = flatten_ids.to(torch.int64)
    m2 = means2d.reshape(C * N, 2)
    cn = conics.reshape(C * N, 3)
    co = colors.reshape(C * N, D)
    op = opacities.reshape(C * N)

    out_rows = [[None] * tw for _ in range(C * th)]
    alpha_rows = [[None] * tw for _ in range(C * th)]
    n_touched = torch.zeros(C * N, dtype=torch.int64)
    last_ids = torch.zeros(C, height, width, dtype=torch.int32)
    P = 0
    for c in range(C):
        bg = backgrounds[c] if backgrounds is not None else None
        for ty in range(th):
            h0, h1 = ty * tile_size, min((ty + 1) * tile_size, height)
            for tx in range(tw):
                w0, w1 = tx * tile_size, min((tx + 1) * tile_size, width)
                t_lin = (c * th + ty) * tw + tx
                s, e = off[t_lin], off[t_lin + 1]
                npx = (h1 - h0) * (w1 - w0)
                if e <= s:
                    col = torch.zeros(npx, D, dtype=dt)
                    if bg is not None:
                        col = col + bg
                    out_rows[c * th + ty][tx] = col.reshape(h1 - h0, w1 - w0, D)
                    alpha_rows[c * th + ty][tx] = torch.zeros(h1 - h0, w1 - w0, dtype=dt)
                    continue
                ids = fid[s:e]
                py, px = torch.meshgrid(
                    torch.arange(h0, h1, dtype=dt) + 0.5,
                    torch.arange(w0, w1, dtype=dt) + 0.5,
                    indexing="ij",
                )
                px, py = px.reshape(-1, 1), py.reshape(-1, 1)
                dx = m2[ids, 0][None, :] - px
                dy = m2[ids, 1][None, :] - py
                a, b, cc = cn[ids, 0][None], cn[ids, 1][None], cn[ids, 2][None]
                sigma = 0.5 * (a * dx * dx + cc * dy * dy) + b * dx * dy
                alpha = torch.clamp(op[ids][None] * torch.exp(-sigma), max=ALPHA_MAX)
                m = (sigma >= 0) & (alpha >= ALPHA_MIN)
                am = torch.where(m, alpha, torch.zeros_like(alpha))
                q = torch.cumprod(1 - am, dim=1)
                keep = m & (q > T_STOP)
                T_before = torch.cat([torch.ones_like(q[:, :1]), q[:, :-1]], dim=1)
                w = am * T_before * keep.to(dt)
                T_final = torch.prod(torch.where(keep, 1 - am, torch.ones_like(am)), dim=1)
                col = w @ co[ids]
                if bg is not None:
                    col = col + T_final[:, None] * bg
                out_rows[c * th + ty][tx] = col.reshape(h1 - h0, w1 - w0, D)
                alpha_rows[c * th + ty][tx] = (1 - T_final).reshape(h1 - h0, w1 - w0)
                with torch.no_grad():
                    vis = keep & (T_before > visibility_min_T)
                    n_touched.index_add_(0, ids, vis.sum(0))
                    if return_stats:
                        G = e - s
                        stop = m & (q <= T_STOP)
                        has_stop = stop.any(1)
                        first_stop = torch.argmax(stop.to(torch.int8), dim=1)
                        P += int(torch.where(has_stop, first_stop + 1, torch.full_like(first_stop, G)).sum())
                        idx = torch.arange(G)[None].expand_as(keep)
                        # absolute index into flatten_ids of the last accumulated pair (0 if none)
                        last = torch.where(keep, idx + s, torch.zeros_like(idx)).amax(1)
                        last_ids[c, h0:h1, w0:w1] = last.reshape(h1 - h0, w1 - w0).to(torch.int32)
    render = torch.cat([torch.cat(r, dim=1) for r in out_rows], dim=0).reshape(C, height, width, D)
    alphas = torch.cat([torch.cat(r, dim=1) for r in alpha_rows], dim=0).reshape(C, height, width, 1)
    n_touched = n_touched.reshape(C, N).to(torch.int32)
    if return_stats:
        return render, alphas, n_touched, {"last_ids": last_ids, "P": P}
    return render, alphas, n_touched


# ----------------------------------------------------------------------------
# a1/a2  the boundary  (reference gslam/rasterization.py:17-41, 44-360)
# ----------------------------------------------------------------------------
@dataclass
class RasterizationOutput:
    per_gaussian_params: ClassVar[tuple] = ("radii", "means2d")
    rgbs: Tensor = None
    alphas: Tensor = None
    depthmaps: Tensor = None
    betas: Tensor = None
    tile_width: int = None
    tile_height: int = None
    tiles_per_gauss: Tensor = None
    isect_ids: Tensor = None
    flatten_ids: Tensor = None
    isect_offsets: Tensor = None
    width: int = None
    height: int = None
    tile_size: int = None
    n_cameras: int = None
    camera_ids: Tensor = None
    gaussian_ids: Tensor = None
    radii: Tensor = None
    means2d: Tensor = None
    depths: Tensor = None
    conics: Tensor = None
    opacities: Tensor = None
    n_touched: Tensor = None


def rasterization(
    means, quats, log_scales, logit_opacities, logit_colors, viewmats, Ks, width, height,
    near_plane=0.01, far_plane=1e10, radius_clip=0.0, eps2d=0.3, packed=False, tile_size=16,
    backgrounds=None, render_mode="RGB", log_uncertainties=None, visibility_min_T=0.5,
    return_stats=False, sh_degree=None,
):
    """Unpacked path of reference gslam/rasterization.py:44-360 (the only one
    its callers use, gslam/map.py:59-74, 88-103)."""
    assert not packed
    assert render_mode in ("RGB", "D", "RGB+D")   # 'ED' modes raise KeyError upstream (:344)
    N, C = means.shape[0], viewmats.shape[0]
    scales, opacities, colors, betas = activations(
        log_scales, logit_opacities, logit_colors, log_uncertainties)
    radii, means2d, depths, conics = fully_fused_projection(
        means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane, radius_clip)
    opacities = opacities.repeat(C, 1)                                    # :187
    if sh_degree is None:
        colors = colors.expand(C, -1, -1)                                 # :225
    else:                                                                 # extension, see sh_colors
        colors = sh_colors(sh_degree, logit_colors, means, viewmats)
    depth_index = betas_index = None
    if render_mode == "RGB+D":
        colors = torch.cat((colors, depths[..., None]), dim=-1)           # :235
        if backgrounds is not None:
            backgrounds = torch.cat([backgrounds, torch.zeros(C, 1, dtype=backgrounds.dtype)], dim=-1)
        depth_index = colors.shape[-1] - 1
    elif render_mode == "D":
        colors = depths[..., None]                                        # :242
        if backgrounds is not None:
            backgrounds = torch.zeros(C, 1, dtype=backgrounds.dtype)
        depth_index = 0
    if log_uncertainties is not None:
        colors = torch.cat((colors, betas.expand(C, -1)[..., None]), dim=-1)   # :250
        if backgrounds is not None:
            backgrounds = torch.cat(
                [backgrounds, torch.full((C, 1), 1.0, dtype=backgrounds.dtype).exp()], dim=-1)  # :253
        betas_index = colors.shape[-1] - 1
    tile_width = math.ceil(width / float(tile_size))
    tile_height = math.ceil(height / float(tile_size))
    tiles_per_gauss, isect_ids, flatten_ids = isect_tiles(
        means2d, radii, depths, tile_size, tile_width, tile_height)
    isect_offsets = isect_offset_encode(isect_ids, C, tile_width, tile_height)
    res = rasterize_to_pixels(
        means2d, conics, colors, opacities, width, height, tile_size, isect_offsets, flatten_ids,
        backgrounds=backgrounds, visibility_min_T=visibility_min_T, return_stats=return_stats)
    render_colors, render_alphas, n_touched = res[:3]
    out = RasterizationOutput(
        alphas=render_alphas, tile_width=tile_width, tile_height=tile_height,
        tiles_per_gauss=tiles_per_gauss, isect_ids=isect_ids, flatten_ids=flatten_ids,
        isect_offsets=isect_offsets, width=width, height=height, tile_size=tile_size,
        n_cameras=C, radii=radii, means2d=means2d, depths=depths, conics=conics,
        opacities=opacities, n_touched=n_touched.long())
    if depth_index is not None:
        out.depthmaps = render_colors[..., depth_index]                   # :341
    if betas_index is not None:
        out.betas = render_colors[..., betas_index]                       # :346
    out.rgbs = render_colors[..., :3] if render_mode != "D" else None     # :347-350
    if return_stats:
        return out, res[3]
    return out


# ----------------------------------------------------------------------------
# a10  pose chain  (reference gslam/primitives.py:15-36, 82-92)
# ----------------------------------------------------------------------------
def rotation_6d_to_matrix(d6: Tensor) -> Tensor:
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = F.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def pose_zhou(Rt: Tensor, dR: Tensor, dt: Tensor) -> Tensor:
    """viewmat = Rt @ [[GS(dR + id6), dt], [0 0 0 1]]  (gslam/primitives.py:82-92)."""
    ident = torch.tensor([1.0, 0, 0, 0, 1.0, 0], dtype=Rt.dtype)
    rot = rotation_6d_to_matrix(dR + ident)
    top = torch.cat([rot, dt.view(3, 1)], dim=-1)
    bottom = torch.tensor([[0, 0, 0, 1.0]], dtype=Rt.dtype)
    return Rt @ torch.cat([top, bottom], dim=0)


# ----------------------------------------------------------------------------
# K9  fused_ssim  (reference call gslam/backend.py:303-307; fused-ssim@30fb258)
# ----------------------------------------------------------------------------
def gaussian_window(size: int = 11, sigma: float = 1.5, dtype=torch.float32) -> Tensor:
    x = torch.arange(size, dtype=torch.float64) - size // 2
    g = torch.exp(-(x * x) / (2 * sigma * sigma))
    return (g / g.sum()).to(dtype)


def ssim_map(img1: Tensor, img2: Tensor) -> Tensor:
    """Per-pixel SSIM of NCHW images: separable 11-tap sigma=1.5 window, zero
    padding ('same'), C1 = 0.01^2, C2 = 0.03^2, each channel on its own."""
    Cn = img1.shape[1]
    g = gaussian_window(dtype=img1.dtype)
    win = (g[:, None] * g[None, :])[None, None].expand(Cn, 1, 11, 11).contiguous()
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    blur = lambda x: F.conv2d(x, win, padding=5, groups=Cn)
    mu1, mu2 = blur(img1), blur(img2)
    s11 = blur(img1 * img1) - mu1 * mu1
    s22 = blur(img2 * img2) - mu2 * mu2
    s12 = blur(img1 * img2) - mu1 * mu2
    return ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s11 + s22 + C2))


def fused_ssim(img1: Tensor, img2: Tensor, padding: str = "same") -> Tensor:
    m = ssim_map(img1, img2)
    if padding == "valid":
        m = m[:, :, 5:-5, 5:-5]
    return m.mean()


# ----------------------------------------------------------------------------
# K10/K11  losses of the mapping and tracking steps
# ----------------------------------------------------------------------------
def exposure_affine(rgbs: Tensor, exposure_params: Tensor) -> Tensor:
    """rgb * e^a + b per view (gslam/backend.py:273-275, gslam/frontend.py:632-636)."""
    return rgbs * exposure_params[..., 0].view(-1, 1, 1, 1).exp() + exposure_params[..., 1].view(-1, 1, 1, 1)


def mapping_photometric_loss(rendered_rgbs: Tensor, gt_imgs: Tensor, betas: Tensor) -> Tensor:
    """active-NeRF NLL (gslam/backend.py:277-283)."""
    l = (rendered_rgbs - gt_imgs).square().sum(dim=-1) / (2 * betas.square())
    return l.mean() + (betas.log().square() * 0.5).mean()


def tracking_loss(gt_img: Tensor, rendered_img: Tensor, betas: Tensor) -> Tensor:
    """'active-nerf' branch of gslam/frontend.py:113-138."""
    err = rendered_img - gt_img
    return (err.square().sum(dim=-1) * betas.pow(-2.0)).mean()


def isotropic_loss(log_scales: Tensor, radii: Tensor) -> Tensor:
    """gslam/backend.py:286-296."""
    visible = radii.sum(dim=0) > 0
    mean_scales = log_scales[visible].mean(dim=1, keepdim=True).exp().detach()
    return (log_scales.exp()[visible] - mean_scales).abs().sum()


def edge_aware_tv(depth: Tensor, rgb: Tensor, mask: Tensor) -> Tensor:
    """gslam/utils.py:136-161."""
    gdx = torch.abs(depth[..., :, :-1, None] - depth[..., :, 1:, None])
    gdy = torch.abs(depth[..., :-1, :, None] - depth[..., 1:, :, None])
    gix = torch.mean(torch.abs(rgb[..., :, :-1, :] - rgb[..., :, 1:, :]), -1, keepdim=True)
    giy = torch.mean(torch.abs(rgb[..., :-1, :, :] - rgb[..., 1:, :, :]), -1, keepdim=True)
    gdx = gdx * torch.exp(-gix)
    gdy = gdy * torch.exp(-giy)
    return gdx[mask[..., :, :-1, None]].sum() + gdy[mask[..., :-1, :, None]].sum()


def mapping_loss(out: RasterizationOutput, gt_imgs, exposure_params, log_scales,
                 ssim_weight=0.2, iso_weight=5e-4, depth_reg_weight=1e-6, regularize=True):
    """total loss of one mapping iteration (gslam/backend.py:273-318)."""
    rendered = exposure_affine(out.rgbs, exposure_params)
    photo = mapping_photometric_loss(rendered, gt_imgs, out.betas)
    iso = isotropic_loss(log_scales, out.radii)
    ssim = 1.0 - fused_ssim(out.rgbs.permute(0, 3, 1, 2), gt_imgs.permute(0, 3, 1, 2), padding="valid")
    total = (1.0 - ssim_weight) * photo + ssim_weight * ssim + iso_weight * iso
    if regularize:
        total = total + depth_reg_weight * edge_aware_tv(out.depthmaps, out.rgbs, out.alphas[..., 0] > 0.4)
    return total, photo


# ----------------------------------------------------------------------------
# K12/K13  Adam (torch.optim.Adam defaults, gslam/backend.py:565-602) + opacity decay (:356-359)
# ----------------------------------------------------------------------------
def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """One Adam update, returns new (param, exp_avg, exp_avg_sq); `step` is the
    1-based count AFTER this update (as torch.optim.Adam)."""
    exp_avg = beta1 * exp_avg + (1 - beta1) * grad
    exp_avg_sq = beta2 * exp_avg_sq + (1 - beta2) * grad * grad
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = exp_avg_sq.sqrt() / math.sqrt(bc2) + eps
    param = param - (lr / bc1) * exp_avg / denom
    return param, exp_avg, exp_avg_sq


def opacity_decay(logit_opacities: Tensor, radii: Tensor, decay: float = 0.995) -> Tensor:
    seen = (radii > 0).sum(dim=0) > 1
    return torch.where(seen, logit_opacities * decay, logit_opacities)
