"""Host-side logic of keyframe insertion and densification (no GPU): the decision rules of
gslam/insertion.py:117-273 and :276-347 are torch expressions that run on any device; only the
append itself (surgery kernel) and the occlusion test (packed projection) need CUDA."""
import math

import torch

from gslam_b200.insertion import (InsertFromDepthMap, InsertUsingImagePlaneGradients, backproject,
                                  knn_mean_distance)
from gslam_b200.rasterization import RasterizationOutput
from gslam_b200.synthetic import intrinsics


def test_knn_mean_distance_on_a_grid():
    # unit grid: every interior point has >= 3 neighbours at distance 1
    g = torch.stack(torch.meshgrid(torch.arange(6.0), torch.arange(6.0), indexing="ij"), -1).reshape(-1, 2)
    pts = torch.cat([g, torch.zeros(36, 1)], dim=1)
    d = knn_mean_distance(pts, 3)
    interior = ((g > 0) & (g < 5)).all(dim=1)
    assert torch.allclose(d[interior], torch.ones(int(interior.sum())))
    corner = d[0]                                   # neighbours at 1, 1, sqrt(2)
    assert abs(float(corner) - (2 + math.sqrt(2)) / 3) < 1e-6
    assert torch.allclose(knn_mean_distance(pts, 3, chunk=7), d)
    assert knn_mean_distance(pts[:1], 3).shape == (1,)


def test_densification_masks_follow_the_reference_rule():
    # gslam/insertion.py:300-317: grads scaled by (W/2 C, H/2 C), norm, mean over views, threshold
    C, N, W, H = 2, 5, 64, 48
    grad = torch.zeros(C, N, 2)
    grad[:, 1, 0] = 1e-3                             # strong, small splat  -> duplicate
    grad[:, 2, 1] = 1e-3                             # strong, large splat  -> split
    grad[0, 3, 0] = 1e-7                             # weak                 -> nothing
    log_scales = torch.full((N, 3), math.log(0.005))
    log_scales[2, 1] = math.log(0.5)
    ins = InsertUsingImagePlaneGradients(0.0002, 0.01)
    dup, split = ins.masks(grad, log_scales, W, H, C)
    assert dup.tolist() == [False, True, False, False, False]
    assert split.tolist() == [False, False, True, False, False]
    g = grad.clone()
    g[..., 0] *= W / 2.0 * C
    g[..., 1] *= H / 2.0 * C
    ref = g.norm(dim=-1).mean(dim=0) > 0.0002
    assert torch.equal(dup | split, ref)


def _mock(alpha_value, depth, H, W):
    return RasterizationOutput(None, torch.full((1, H, W, 1), alpha_value), depth.clone().view(1, H, W))


def test_propose_counts_depths_and_geometry_without_a_map():
    K, W, H = intrinsics("tum", 0.25)
    g = torch.Generator().manual_seed(0)
    depth = 1.0 + 0.1 * torch.randn(H, W, generator=g)
    img = torch.rand(H, W, 3, generator=g)
    ins = InsertFromDepthMap(0.1, 0.2, 0.1, 0.3, False)
    # nothing covered (alpha 0.01 < 0.1): all N go to open pixels at median depth + noise (:147-165)
    new = ins.propose(_mock(0.01, depth, H, W), K, torch.eye(4), img, None, 300, generator=g)
    assert new["means"].shape == (300, 3) and new["quats"].shape == (300, 4)
    assert torch.allclose(new["opacities"], torch.full((300,), math.log(0.3 / 0.7)), atol=1e-6)
    assert torch.equal(new["log_uncertainties"], torch.ones(300))
    z = new["means"][:, 2]
    assert float(z.min()) >= 0.1 and abs(float(z.mean()) - float(depth.median())) < 0.1
    u = new["means"][:, 0] / z * K[0, 0] + K[0, 2]
    v = new["means"][:, 1] / z * K[1, 1] + K[1, 2]
    assert bool(((u > -0.5) & (u < W - 0.5) & (v > -0.5) & (v < H - 0.5)).all())
    # no map: per-splat scales are the 3-NN mean distances (:226-231)
    assert torch.allclose(torch.exp(new["scales"][:, 0]), knn_mean_distance(new["means"], 3), rtol=1e-5)
    # colours are the logit of the image at the picked pixels: sigmoid brings them back into (0, 1)
    assert bool((torch.sigmoid(new["colors"]) > 0).all()) and bool((torch.sigmoid(new["colors"]) < 1).all())


def test_propose_split_between_open_and_covered_pixels():
    K, W, H = intrinsics("tum", 0.25)
    g = torch.Generator().manual_seed(1)
    depth = torch.full((H, W), 2.0)
    img = torch.rand(H, W, 3, generator=g)
    out = _mock(1.0, depth, H, W)                    # everything covered ...
    out.alphas[0, :4] = 0.0                          # ... except four rows: 4 W open pixels
    n_open = 4 * W
    for with_depth, N in ((True, 1000), (False, 1000), (True, 100)):
        ins = InsertFromDepthMap(0.001, 0.002, 0.1, 0.3, insert_in_regions_with_depth=with_depth)
        new = ins.propose(out, K, torch.eye(4), img, torch.full((3,), 0.02), N, generator=g)
        in_open = min(N, n_open)
        in_cov = max(0, min(N // 2 - in_open, H * W - n_open)) if with_depth else 0
        assert new["means"].shape[0] == in_open + in_cov
        assert torch.allclose(torch.exp(new["scales"]), torch.full((in_open + in_cov, 3), 0.02))
        assert abs(float(new["means"][:, 2].median()) - 2.0) < 0.01


def test_backproject_matches_the_pinhole_model():
    K, W, H = intrinsics("tum", 0.1)
    depth = torch.full((H, W), 3.0)
    pts = backproject(K, depth).view(H, W, 3)
    assert torch.allclose(pts[..., 2], depth)
    assert abs(float(pts[10, 20, 0]) - (20 - float(K[0, 2])) * 3.0 / float(K[0, 0])) < 1e-5
    assert abs(float(pts[10, 20, 1]) - (10 - float(K[1, 2])) * 3.0 / float(K[1, 1])) < 1e-5
