"""The pruning criteria (gslam/pruning.py:61-139) are [N] torch expressions: checked here on the CPU
against hand-worked cases; the removal itself (surgery kernel) is covered by the GPU tests."""
import math

import torch

from gslam_b200.pruning import (ill_conditioned_mask, large_footprint_mask, low_opacity_mask,
                                young_and_unseen_mask)


def test_low_opacity_mask():
    op = torch.tensor([0.05, 0.19, 0.21, 0.45, 0.9])
    logits = torch.log(op / (1 - op))
    assert low_opacity_mask(logits, 0.2).tolist() == [True, True, False, False, False]
    assert low_opacity_mask(logits, 0.5).tolist() == [True, True, True, True, False]


def test_large_footprint_mask_takes_the_maximum_over_views():
    radii = torch.tensor([[10, 300, 0, 256], [257, 20, 0, 256]], dtype=torch.int32)
    assert large_footprint_mask(radii, 256).tolist() == [True, True, False, False]
    assert large_footprint_mask(radii.max(dim=0).values, 256).tolist() == [True, True, False, False]


def test_ill_conditioned_mask_counts_projected_but_untouched_views():
    radii = torch.tensor([[5, 5, 0, 5]] * 5, dtype=torch.int32)            # 5 views
    n_touched = torch.tensor([[0, 3, 0, 0]] * 4 + [[0, 3, 0, 9]])
    # splat 0: projected, never touched in 5 views (> 3); splat 3: untouched in 4 views (> 3); splat 2 never projected
    assert ill_conditioned_mask(radii, n_touched, 3).tolist() == [True, False, False, True]
    assert ill_conditioned_mask(radii, n_touched, 4).tolist() == [True, False, False, False]


def test_young_and_unseen_mask():
    ages = torch.tensor([0, 7, 8, 9, 9])
    vis = torch.tensor([0, 0, 1, 5, 2])
    # latest keyframe age 9, window 3: ages > 6 are young; fewer than 3 sightings go
    assert young_and_unseen_mask(ages, vis, 9, 3).tolist() == [False, True, True, False, True]
