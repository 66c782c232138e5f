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


def _reference_proposal(depthmap, alphas, K, pose, img, splat_scales, N, min_alpha, depth_var, no_depth_var,
                        initial_opacity, insert_in_regions_with_depth, gt_depthmap=None):
    """A literal restatement of the proposal half of InsertFromDepthMap.step, reference
    gslam/insertion.py:128-243 (statement order, in-place updates and RNG draws as there: randn_like, randint
    over the open pixels, randint over the covered pixels, rand for the quaternions), with the Camera /
    Frame accessors inlined (gslam/primitives.py:369-395).  Test oracle for gslam_b200.insertion.propose."""
    depths = (depthmap if gt_depthmap is None else gt_depthmap).clone()
    valid = torch.logical_and(alphas > min_alpha, depths > 0)
    n_valid = valid.sum().item()
    n_invalid = depths.shape.numel() - n_valid
    n_invalid_splats = min(N, n_invalid)
    n_valid_splats = max(0, min(int(N / 2) - n_invalid_splats, n_valid))
    if n_invalid_splats <= 0 and (not insert_in_regions_with_depth and n_valid_splats <= 0):
        return None
    median_depth = depths[valid].median() if valid.any() else depths.median()
    random_depths = torch.randn_like(depths)
    depths[valid] += random_depths[valid] * depth_var
    depths[~valid] = median_depth
    depths[~valid] += random_depths[~valid] * no_depth_var
    depths.clamp_min_(0.1)
    idx_valid = torch.nonzero(valid.reshape(-1))
    idx_invalid = torch.nonzero(~valid.reshape(-1))
    picks = []
    if n_invalid_splats > 0:
        picks.append(idx_invalid[torch.randint(idx_invalid.shape[0], [n_invalid_splats])].reshape(-1))
    if insert_in_regions_with_depth and n_valid_splats > 0:
        picks.append(idx_valid[torch.randint(idx_valid.shape[0], [n_valid_splats])].reshape(-1))
    if len(picks) == 0:
        return None
    picks = torch.cat(picks)
    n = picks.shape[0]
    H, W = depths.shape
    v, u = torch.meshgrid(torch.arange(H, dtype=depths.dtype), torch.arange(W, dtype=depths.dtype), indexing="ij")
    pts = torch.stack([(u - K[0, 2]) * depths / K[0, 0], (v - K[1, 2]) * depths / K[1, 1], depths], -1).reshape(-1, 3)
    means = pts[picks]
    colors = img.reshape([-1, 3])[picks]
    c2w = torch.linalg.inv(pose)
    means = means @ c2w[:3, :3].t() + c2w[:3, 3]
    scales = torch.exp(splat_scales).median(dim=0)[0].tile([n, 1])
    return {"means": means, "scales": torch.log(scales), "colors": torch.logit(colors),
            "opacities": torch.logit(torch.full((n,), initial_opacity)), "quats": torch.rand((n, 4)),
            "log_uncertainties": torch.ones((n,))}


def test_proposal_matches_a_literal_restatement_of_the_reference_draw_for_draw():
    """Same seed, same draws, same candidates: the proposal of gslam_b200.insertion against the
    reference's statement sequence, for a mixed covered / open frame, an all-open frame, and RGB-D input."""
    K, W, H = intrinsics("tum", 0.1)
    g = torch.Generator().manual_seed(5)
    pose = torch.eye(4)
    pose[:3, 3] = torch.tensor([0.05, -0.02, 0.1])
    ang = 0.1
    pose[:3, :3] = torch.tensor([[math.cos(ang), 0, math.sin(ang)], [0, 1, 0], [-math.sin(ang), 0, math.cos(ang)]])
    img = torch.rand(H, W, 3, generator=g) * 0.98 + 0.01
    splat_scales = torch.log(0.01 + 0.02 * torch.rand(500, 3, generator=g))
    depthmap = 1.0 + torch.rand(H, W, generator=g)
    gt_depth = 2.0 + torch.rand(H, W, generator=g)
    gt_depth[:3] = 0.0
    cases = []
    a_mixed = torch.rand(H, W, generator=g)                    # ~90 % covered at min_alpha 0.1
    cases.append(("mixed", a_mixed, 40, True, None))
    cases.append(("mostly covered, covered region allowed", (a_mixed > 0.002).float(), 300, True, None))
    cases.append(("covered region not allowed", a_mixed, 40, False, None))
    cases.append(("all open", torch.zeros(H, W), 25, True, None))
    cases.append(("rgbd", a_mixed, 60, True, gt_depth))
    for name, alphas, N, allow, gtd in cases:
        ins = InsertFromDepthMap(0.1, 0.2, 0.1, 0.3, allow)
        out = RasterizationOutput(None, alphas[None, ..., None].clone(), depthmap[None].clone())
        torch.manual_seed(1234)
        ref = _reference_proposal(depthmap, alphas, K, pose, img, splat_scales, N, 0.1, 0.1, 0.2, 0.3, allow, gtd)
        torch.manual_seed(1234)
        ours = ins.propose(out, K, pose, img, torch.exp(splat_scales).median(dim=0)[0], N, gtd, generator=None)
        assert (ref is None) == (ours is None), name
        if ref is None:
            continue
        assert set(ref) == set(ours), name
        for k in ref:
            assert ref[k].shape == ours[k].shape, (name, k)
            torch.testing.assert_close(ours[k], ref[k], rtol=1e-5, atol=1e-6, msg=f"{name}: {k}")
