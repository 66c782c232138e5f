"""Pruning criteria of the mapping loop on the flat splat layout.

The masks follow reference gslam/pruning.py:61-139 (low opacity, large screen footprint, visible
but touching no pixel in too many views, young and rarely visible) and their combination at the end
of `optimize_map` (gslam/backend.py:364-391); the removal itself is `surgery.prune_using_mask`, one
launch over the parameters and the Adam state.  The criteria are a handful of [N] elementwise
expressions evaluated once per mapping round, so they stay torch expressions on the device.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor

from .mapping import FlatSplats
from .surgery import prune_using_mask


def low_opacity_mask(logit_opacities: Tensor, min_opacity: float) -> Tensor:
    """PruneLowOpacity (gslam/pruning.py:61-73)."""
    return torch.sigmoid(logit_opacities) < min_opacity


def large_footprint_mask(radii: Tensor, max_radius: float) -> Tensor:
    """PruneLargeGaussians (:98-114); radii [C,N] or the per-Gaussian maximum [N]."""
    per_gauss = radii.max(dim=0).values if radii.dim() == 2 else radii
    return per_gauss > max_radius


def ill_conditioned_mask(radii: Tensor, n_touched: Tensor, max_views: int) -> Tensor:
    """PruneIllConditionedGaussians (:117-139): projected (radius > 0) yet touching no pixel in more
    than max_views of the [C,N] views."""
    return ((radii > 0) & (n_touched == 0)).sum(dim=0) > max_views


def young_and_unseen_mask(ages: Tensor, visibility_counts: Tensor, latest_kf_age: int, min_visibility: int,
                          young_window: int = 3) -> Tensor:
    """PruneByVisibility (:76-95): inserted within the last `young_window` keyframes and seen fewer
    than min_visibility times."""
    return (ages > (latest_kf_age - young_window)) & (visibility_counts < min_visibility)


@torch.no_grad()
def prune_remove_mask(logit_opacities: Tensor, radii: Tensor, n_touched: Tensor, min_opacity: float,
                      max_radius: float, max_unproductive_views: Optional[int] = None, window_last_n: int = 8,
                      group=None, n_views_total: Optional[int] = None) -> Tensor:
    """The remove mask of the pruning block at the end of optimize_map (gslam/backend.py:364-391): OR of the
    size and opacity criteria and, when max_unproductive_views is given and the window has >= 2 views, the
    conditioning criterion.  In a view-sharded step `radii` / `n_touched` hold only THIS rank's views: the
    per-Gaussian maximum radius is MAX-all-reduced and the count of unproductive views SUM-all-reduced over
    `group`, so every rank derives the same mask and performs the same surgery."""
    from .mapping import allreduce_max_, allreduce_sum_
    radii_max = allreduce_max_(radii.max(dim=0).values.contiguous(), group) if radii.shape[0] else \
        allreduce_max_(radii.new_zeros(radii.shape[1]), group)
    remove = (radii_max > max_radius) | low_opacity_mask(logit_opacities, min_opacity)
    n_views = radii.shape[0] if n_views_total is None else n_views_total
    if max_unproductive_views is not None and n_views >= 2:
        r, t = (radii, n_touched) if n_views_total is not None else (radii[:window_last_n], n_touched[:window_last_n])
        unproductive = allreduce_sum_(((r > 0) & (t == 0)).sum(dim=0).contiguous(), group)
        remove |= unproductive > max_unproductive_views
    return remove


@torch.no_grad()
def prune_round(splats: FlatSplats, optimizer, radii: Tensor, n_touched: Tensor, min_opacity: float,
                max_radius: float, max_unproductive_views: Optional[int] = None, window_last_n: int = 8,
                per_gaussian_params: Optional[list] = None, group=None, n_views_total: Optional[int] = None) -> int:
    """`prune_remove_mask` + the surgery (one launch over the parameters and the Adam state); returns the
    number removed."""
    remove = prune_remove_mask(splats.params[3].detach(), radii, n_touched, min_opacity, max_radius,
                               max_unproductive_views, window_last_n, group, n_views_total)
    return prune_using_mask(splats, optimizer, ~remove, per_gaussian_params=per_gaussian_params)
