"""Keyframe insertion from a depth map, on the flat splat layout.

Follows the algorithm of reference `InsertFromDepthMap.step` (gslam/insertion.py:117-273):

1. pixels with rendered alpha above a threshold and positive depth count as "covered"; up to N new
   splats go to uncovered pixels first, and (optionally) up to N/2 minus that many to covered ones;
2. the depth of a new splat is the rendered depth plus noise where covered, the median covered
   depth plus (larger) noise elsewhere, clamped to 0.1;
3. the chosen pixels are back-projected with the frame's intrinsics and moved to world space with
   the inverse of the frame's view matrix; colour = logit of the image, scale = the map's median
   scale, quaternion random, opacity / uncertainty at their initial values;
4. a candidate that lies IN FRONT of the estimated depth of any earlier keyframe at the pixel it
   projects to is dropped (packed projection, `get_new_splat_depth`, gslam/rasterization.py:363-448);
5. the survivors are appended to the parameters and to the Adam state in one launch
   (`surgery.add_new_splats`, replacing `_add_new_splats`, gslam/insertion.py:27-62).

Every tensor operation runs on the device; the two host reads are the covered-pixel count (as in
the reference, :137) and the number of survivors that sizes the new buffers.  Random choices take an
optional `torch.Generator` so that tests are repeatable.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import torch
from torch import Tensor

from .mapping import FlatSplats
from .primitives import affine_inverse
from .rasterization import RasterizationOutput, get_new_splat_depth
from .surgery import add_new_splats


@dataclass
class KeyframeView:
    """What step 4 needs from an earlier keyframe (Frame.camera.intrinsics, Frame.pose(), Frame.est_depths)."""
    intrinsics: Tensor      # [3,3]
    viewmat: Tensor         # [4,4] world -> camera
    est_depths: Tensor      # [H,W]


def backproject(intrinsics: Tensor, depth_map: Tensor) -> Tensor:
    """Camera-space points [H*W, 3] of a depth map (Camera.backproject, gslam/primitives.py:369-395)."""
    H, W = depth_map.shape
    fx, fy, cx, cy = intrinsics[0, 0], intrinsics[1, 1], intrinsics[0, 2], intrinsics[1, 2]
    v, u = torch.meshgrid(torch.arange(H, device=depth_map.device, dtype=depth_map.dtype),
                          torch.arange(W, device=depth_map.device, dtype=depth_map.dtype), indexing="ij")
    x = (u - cx) * (depth_map / fx)
    y = (v - cy) * (depth_map / fy)
    return torch.stack([x, y, depth_map], dim=-1).reshape(-1, 3)


def knn_mean_distance(points: Tensor, k: int = 3, chunk: int = 4096) -> Tensor:
    """Mean distance of every point [n,3] to its k nearest neighbours (the `knn(means, 4)[:, 1:]` of
    gslam/insertion.py:226-230); dense distances in row chunks -- n is at most a few thousand here."""
    n = points.shape[0]
    if n <= 1:
        return torch.ones(n, device=points.device, dtype=points.dtype)
    kk = min(k, n - 1)
    out = []
    for i in range(0, n, chunk):
        # exact differences, not the |a|^2 + |b|^2 - 2ab GEMM form (cancellation at small distances)
        d = torch.cdist(points[i:i + chunk], points, compute_mode="donot_use_mm_for_euclid_dist")
        out.append(d.topk(kk + 1, dim=1, largest=False).values[:, 1:].mean(dim=1))
    return torch.cat(out)


class InsertFromDepthMap:
    def __init__(self, depth_variance: float, no_depth_variance: float, min_alpha_for_depth: float,
                 initial_opacity: float, insert_in_regions_with_depth: bool = True):
        self.depth_variance = depth_variance
        self.no_depth_variance = no_depth_variance
        self.min_alpha_for_depth = min_alpha_for_depth
        self.initial_opacity = initial_opacity
        self.insert_in_regions_with_depth = insert_in_regions_with_depth

    @torch.no_grad()
    def propose(self, rasterization_output: RasterizationOutput, intrinsics: Tensor, viewmat: Tensor,
                img: Tensor, median_scale: Optional[Tensor], N: int, gt_depthmap: Optional[Tensor] = None,
                generator: Optional[torch.Generator] = None) -> Optional[Dict[str, Tensor]]:
        """Steps 1-3: the candidate splats (parameter dict in the map's activation space) or None."""
        depths = (rasterization_output.depthmaps[0] if gt_depthmap is None else gt_depthmap).clone()
        alphas = rasterization_output.alphas[0, ..., 0]
        dev = depths.device
        covered = (alphas > self.min_alpha_for_depth) & (depths > 0)
        n_covered = int(covered.sum().item())
        n_open = depths.numel() - n_covered
        n_in_open = min(N, n_open)
        n_in_covered = max(0, min(N // 2 - n_in_open, n_covered))
        if n_in_open <= 0 and not (self.insert_in_regions_with_depth and n_in_covered > 0):
            return None
        median_depth = depths[covered].median() if n_covered else depths.median()
        noise = torch.randn(depths.shape, device=dev, dtype=depths.dtype, generator=generator)
        depths = torch.where(covered, depths + noise * self.depth_variance,
                             median_depth + noise * self.no_depth_variance).clamp_min_(0.1)
        flat_cov = covered.reshape(-1)
        picks: List[Tensor] = []
        for mask, count in ((~flat_cov, n_in_open),
                            (flat_cov, n_in_covered if self.insert_in_regions_with_depth else 0)):
            if count > 0:
                pool = torch.nonzero(mask).reshape(-1)
                picks.append(pool[torch.randint(pool.shape[0], (count,), device=dev, generator=generator)])
        if not picks:
            return None
        pick = torch.cat(picks)
        n = pick.shape[0]
        cam_pts = backproject(intrinsics, depths)[pick]
        c2w = affine_inverse(viewmat)          # = torch.linalg.inv for a view matrix, without the LU + sync
        means = cam_pts @ c2w[:3, :3].t() + c2w[:3, 3]
        colors = img.reshape(-1, 3)[pick].clamp(1e-4, 1 - 1e-4)      # logit of exact 0 / 1 is infinite
        if median_scale is None:
            # no map yet: each splat's scale is its mean distance to its 3 nearest neighbours (:226-231)
            scales = knn_mean_distance(means, 3).clamp_min(1e-6).unsqueeze(-1).repeat(1, 3)
        else:
            scales = median_scale.expand(n, 3)
        return {
            "means": means,
            "scales": torch.log(scales).contiguous(),
            "colors": torch.logit(colors),
            "opacities": torch.logit(torch.full((n,), self.initial_opacity, device=dev)),
            "quats": torch.rand(n, 4, device=dev, generator=generator),
            "log_uncertainties": torch.ones(n, device=dev),
        }

    @torch.no_grad()
    def occlusion_filter(self, new_params: Dict[str, Tensor], keyframes: Sequence[KeyframeView], width: int,
                         height: int) -> Tensor:
        """Step 4: keep mask [n]; False where a candidate is in front of an earlier keyframe's depth."""
        n = new_params["means"].shape[0]
        keep = torch.ones(n, dtype=torch.bool, device=new_params["means"].device)
        if len(keyframes) <= 1 or n == 0:
            return keep
        Ks = torch.stack([k.intrinsics for k in keyframes])
        viewmats = torch.stack([k.viewmat for k in keyframes])
        est = torch.stack([k.est_depths for k in keyframes])
        camera_ids, gaussian_ids, _, means2d, depths = get_new_splat_depth(new_params, viewmats, Ks, width, height)
        px = means2d.to(torch.long)
        col = px[:, 0].clamp(0, width - 1)
        row = px[:, 1].clamp(0, height - 1)
        in_front = depths < est[camera_ids, row, col]
        keep[gaussian_ids[in_front]] = False
        return keep

    @torch.no_grad()
    def step(self, splats: FlatSplats, optimizer, rasterization_output: RasterizationOutput,
             intrinsics: Tensor, viewmat: Tensor, img: Tensor, N: int,
             keyframes: Sequence[KeyframeView] = (), gt_depthmap: Optional[Tensor] = None,
             generator: Optional[torch.Generator] = None) -> int:
        """Insert up to N splats seen from (intrinsics, viewmat); returns how many were added."""
        median_scale = torch.exp(splats.params[2].detach()).median(dim=0)[0] if splats.n else None
        new_params = self.propose(rasterization_output, intrinsics, viewmat, img, median_scale, N,
                                  gt_depthmap, generator)
        if new_params is None:
            return 0
        H, W = rasterization_output.alphas.shape[1:3]
        keep = self.occlusion_filter(new_params, keyframes, W, H)
        if not bool(keep.all()):
            new_params = {k: v[keep] for k, v in new_params.items()}
        return add_new_splats(splats, optimizer, new_params)


class InsertUsingImagePlaneGradients:
    """The 3DGS densification rule of reference gslam/insertion.py:276-347 on the flat layout.

    A splat whose image-plane gradient norm (averaged over the window's views, scaled by half the
    image size times the view count, :300-309) exceeds `grow_grad2d` is DUPLICATED when its largest
    scale is at most `grow_scale3d` and SPLIT otherwise: the split copy is displaced by
    `covar @ N(0, I)` and shrunk by 1/1.6 (:80-97).  Copies start with log-uncertainty 1 and zero
    Adam moments; the originals stay (as in the reference).  Both sets are appended in ONE surgery
    launch (the reference runs `_add_new_splats` twice, each a `cat` per tensor and per moment).
    """

    def __init__(self, grow_grad2d: float, grow_scale3d: float):
        self.grow_grad2d = grow_grad2d
        self.grow_scale3d = grow_scale3d

    @torch.no_grad()
    def masks(self, means2d_grad: Tensor, log_scales: Tensor, width: int, height: int, n_cameras: int, group=None):
        """(to_duplicate, to_split) [N] from means2d.grad [C,N,2].  In a view-sharded step means2d_grad holds
        this rank's views and n_cameras is the window's total: the per-Gaussian sum of the gradient norms is
        all-reduced over `group`, so every rank derives the same masks."""
        from .mapping import allreduce_sum_
        g = means2d_grad * means2d_grad.new_tensor([width / 2.0 * n_cameras, height / 2.0 * n_cameras])
        high = allreduce_sum_(g.norm(dim=-1).sum(dim=0).contiguous(), group) / max(n_cameras, 1) > self.grow_grad2d
        small = torch.exp(log_scales).max(dim=-1).values <= self.grow_scale3d
        return high & small, high & ~small

    @torch.no_grad()
    def step(self, splats: FlatSplats, optimizer, rasterization_output: RasterizationOutput,
             generator: Optional[torch.Generator] = None, group="local", n_views_total: Optional[int] = None):
        """Returns (num_duplicate, num_split); `rasterization_output.means2d.grad` must be populated
        (`out.means2d.retain_grad()` before backward, gslam/backend.py:326).  group / n_views_total: the
        view-sharded step (a rank without a view passes an output whose means2d is None)."""
        from .rasterization import quat_scale_to_covar_preci
        from .mapping import PARAM_NAMES
        out = rasterization_output
        P = {name: p.detach() for name, p in zip(PARAM_NAMES, splats.params)}
        if out.means2d is None:
            grad = P["means"].new_zeros(0, splats.n, 2)
        else:
            grad = out.means2d.grad
            if grad is None:
                raise RuntimeError("InsertUsingImagePlaneGradients: means2d.grad is empty (call retain_grad())")
        n_cam = out.n_cameras if n_views_total is None else n_views_total
        dup, split = self.masks(grad, P["scales"], out.width, out.height, n_cam, group)
        both = torch.cat([torch.nonzero(dup).reshape(-1), torch.nonzero(split).reshape(-1)])
        n_dup = int(dup.sum().item())                      # the reference reads both counts too (:318-319)
        n_split = both.numel() - n_dup
        if both.numel() == 0:
            return 0, 0
        new = {name: p[both].clone() for name, p in P.items()}
        if n_split:
            covars, _ = quat_scale_to_covar_preci(new["quats"][n_dup:], torch.exp(new["scales"][n_dup:]),
                                                  compute_preci=False)
            noise = torch.randn(n_split, 3, device=covars.device, generator=generator)
            new["means"][n_dup:] += torch.einsum("bij,bj->bi", covars, noise)
            new["scales"][n_dup:] -= math.log(1.6)
        new["log_uncertainties"].fill_(1.0)
        add_new_splats(splats, optimizer, new)
        if out.radii is not None:
            out.radii = torch.cat([out.radii, out.radii.new_zeros(out.radii.shape[0], both.numel())], dim=1)
        return n_dup, n_split
