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
def prune_round(splats: FlatSplats, optimizer, radii: Tensor, n_touched: Tensor, min_opacity: float,
                max_radius: float, max_unproductive_views: Optional[int] = None, window_last_n: int = 8,
                per_gaussian_params: Optional[list] = None) -> int:
    """The pruning block at the end of optimize_map (gslam/backend.py:364-391): OR of the size and opacity
    criteria and, when max_unproductive_views is given and the window has >= 2 views, the
    conditioning criterion over the last `window_last_n` views; returns the number removed."""
    remove = large_footprint_mask(radii, max_radius) | low_opacity_mask(splats.params[3].detach(), min_opacity)
    if max_unproductive_views is not None and radii.shape[0] >= 2:
        remove |= ill_conditioned_mask(radii[:window_last_n], n_touched[:window_last_n], max_unproductive_views)
    return prune_using_mask(splats, optimizer, ~remove, per_gaussian_params=per_gaussian_params)
