"""Keyframe insertion from a depth map: properties of the algorithm of
gslam/insertion.py:117-273 on the flat splat layout (random choices make a bitwise oracle moot)."""
import math

import pytest
import torch

from gslam_b200.insertion import InsertFromDepthMap, KeyframeView, backproject
from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats
from gslam_b200.optim import FusedAdam
from gslam_b200.rasterization import rasterization
from gslam_b200.synthetic import make_scene, target_images

pytestmark = pytest.mark.gpu


def _world(cuda, n=20_000, views=3):
    sc = make_scene(n, views, "tum", 0.5).to(cuda)
    splats = FlatSplats(sc.params())
    opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(splats.params, PARAM_NAMES)])
    for p in splats.params:
        p.grad = torch.zeros_like(p)
    opt.step()                                           # creates the Adam state
    with torch.no_grad():
        out = rasterization(*splats.params[:5], sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                            render_mode="RGB+D", log_uncertainties=splats.params[5],
                            backgrounds=torch.zeros(views, 3, device=cuda))
    return sc, splats, opt, out


def test_backproject_roundtrip(cuda):
    sc = make_scene(10, 1, "tum", 0.5).to(cuda)
    depth = 0.5 + torch.rand(sc.height, sc.width, device=cuda) * 4
    pts = backproject(sc.Ks[0], depth)
    u = pts[:, 0] / pts[:, 2] * sc.Ks[0, 0, 0] + sc.Ks[0, 0, 2]
    v = pts[:, 1] / pts[:, 2] * sc.Ks[0, 1, 1] + sc.Ks[0, 1, 2]
    vv, uu = torch.meshgrid(torch.arange(sc.height, device=cuda), torch.arange(sc.width, device=cuda), indexing="ij")
    assert torch.allclose(u, uu.reshape(-1).float(), atol=1e-3) and torch.allclose(v, vv.reshape(-1).float(), atol=1e-3)
    assert torch.equal(pts[:, 2], depth.reshape(-1))


@pytest.mark.parametrize("with_depth_regions", [True, False])
def test_insert_counts_state_and_geometry(cuda, with_depth_regions):
    sc, splats, opt, out = _world(cuda, n=1200)          # sparse: covered and open pixels both exist
    n0 = splats.n
    ins = InsertFromDepthMap(0.001, 0.002, 0.1, 0.3, insert_in_regions_with_depth=with_depth_regions)
    img = target_images(1, sc.height, sc.width)[0].to(cuda)
    gen = torch.Generator(device=cuda).manual_seed(7)
    first = type(out)(alphas=out.alphas[:1], depthmaps=out.depthmaps[:1])
    covered = int(((out.alphas[0, ..., 0] > 0.1) & (out.depthmaps[0] > 0)).sum())
    n_open = sc.height * sc.width - covered
    N = 400
    assert covered > 1000 and n_open > 1000
    expect = min(N, n_open) + (max(0, min(N // 2 - min(N, n_open), covered)) if with_depth_regions else 0)
    added = ins.step(splats, opt, first, sc.Ks[0], sc.viewmats[0], img, N, generator=gen)
    assert added == expect and splats.n == n0 + added
    new = [p.detach()[n0:] for p in splats.params]
    assert torch.allclose(new[3], torch.full((added,), math.log(0.3 / 0.7), device=cuda), atol=1e-6)
    assert torch.equal(new[5], torch.ones(added, device=cuda))
    assert bool(torch.isfinite(new[4]).all()) and new[1].shape == (added, 4)
    med = torch.exp(splats.params[2].detach()[:n0]).median(dim=0)[0]
    assert torch.allclose(torch.exp(new[2]), med.expand(added, 3), rtol=1e-5)
    # the new centres project back into the frame they were sampled from, at depth >= 0.1
    vm, K = sc.viewmats[0], sc.Ks[0]
    cam = new[0] @ vm[:3, :3].t() + vm[:3, 3]
    u = cam[:, 0] / cam[:, 2] * K[0, 0] + K[0, 2]
    v = cam[:, 1] / cam[:, 2] * K[1, 1] + K[1, 2]
    assert float(cam[:, 2].min()) >= 0.1 - 1e-5
    assert bool(((u > -0.5) & (u < sc.width - 0.5) & (v > -0.5) & (v < sc.height - 0.5)).all())
    for p in splats.params:                               # fresh Adam moments for the appended rows
        st = opt.state[p]
        assert bool((st["exp_avg"][n0:] == 0).all()) and bool((st["exp_avg_sq"][n0:] == 0).all())


def test_occlusion_filter_against_earlier_keyframes(cuda):
    sc, splats, opt, out = _world(cuda)
    ins = InsertFromDepthMap(0.001, 0.002, 0.1, 0.3, True)
    img = target_images(1, sc.height, sc.width)[0].to(cuda)
    gen = torch.Generator(device=cuda).manual_seed(11)
    first = type(out)(alphas=out.alphas[:1], depthmaps=out.depthmaps[:1])
    med = torch.exp(splats.params[2].detach()).median(dim=0)[0]
    cand = ins.propose(first, sc.Ks[0], sc.viewmats[0], img, med, 300, generator=gen)
    n = cand["means"].shape[0]
    H, W = sc.height, sc.width
    near = [KeyframeView(sc.Ks[c], sc.viewmats[c], torch.zeros(H, W, device=cuda)) for c in range(3)]
    far = [KeyframeView(sc.Ks[c], sc.viewmats[c], torch.full((H, W), 1e6, device=cuda)) for c in range(3)]
    assert bool(ins.occlusion_filter(cand, near, W, H).all())          # nothing is in front of depth 0
    keep_far = ins.occlusion_filter(cand, far, W, H)                   # everything visible is in front of 1e6
    assert int(keep_far.sum()) < n // 4
    assert bool(ins.occlusion_filter(cand, far[:1], W, H).all())       # a single keyframe: no test (reference :252)
    # mixed: only the candidates the second keyframe sees in front of its depth 1.0 go
    mixed = [near[0], KeyframeView(sc.Ks[1], sc.viewmats[1], torch.full((H, W), 1.0, device=cuda)), near[2]]
    keep = ins.occlusion_filter(cand, mixed, W, H)
    vm = sc.viewmats[1]
    z1 = (cand["means"] @ vm[:3, :3].t() + vm[:3, 3])[:, 2]
    assert bool((z1[~keep] < 1.0).all())


def test_first_insertion_into_an_empty_map_uses_neighbour_distances(cuda):
    # Backend.initialize (gslam/backend.py:604-631): no map yet, mock depth, 3-NN scales (:226-231)
    from gslam_b200.insertion import knn_mean_distance
    sc = make_scene(10, 1, "tum", 0.5).to(cuda)
    empty = [torch.zeros((0, w) if w > 1 else (0,), device=cuda) for w in (3, 4, 3, 1, 3, 1)]
    splats = FlatSplats(empty)
    opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(splats.params, PARAM_NAMES)])
    H, W = sc.height, sc.width
    from gslam_b200.rasterization import RasterizationOutput
    mock = RasterizationOutput(None, torch.full((1, H, W, 1), 0.01, device=cuda),
                               1.0 + 0.1 * torch.randn(1, H, W, device=cuda))
    img = target_images(1, H, W)[0].to(cuda)
    ins = InsertFromDepthMap(0.1, 0.2, 0.1, 0.3, False)
    added = ins.step(splats, opt, mock, sc.Ks[0], torch.eye(4, device=cuda), img, 500,
                     generator=torch.Generator(device=cuda).manual_seed(3))
    assert added == 500 == splats.n
    means, scales = splats.params[0].detach(), torch.exp(splats.params[2].detach())
    d = torch.cdist(means, means, compute_mode="donot_use_mm_for_euclid_dist")
    ref = d.topk(4, largest=False).values[:, 1:].mean(dim=1)
    assert torch.allclose(knn_mean_distance(means, 3), ref, rtol=1e-5, atol=1e-7)
    assert torch.allclose(scales, ref.clamp_min(1e-6)[:, None].expand(500, 3), rtol=1e-4)
    assert knn_mean_distance(means, 3, chunk=128).allclose(ref, rtol=1e-5, atol=1e-7)


def test_image_plane_gradient_densification(cuda):
    # InsertUsingImagePlaneGradients (gslam/insertion.py:276-347): duplicate small, split large
    from gslam_b200.insertion import InsertUsingImagePlaneGradients
    sc, splats, opt, _ = _world(cuda, n=5000, views=2)
    n0 = splats.n
    P = splats.params
    out = rasterization(*P[:5], sc.viewmats, sc.Ks, sc.width, sc.height, packed=False, render_mode="RGB+D",
                        log_uncertainties=P[5], backgrounds=torch.zeros(2, 3, device=cuda))
    out.means2d.retain_grad()
    gt = target_images(2, sc.height, sc.width).to(cuda)
    (out.rgbs - gt).square().mean().backward()
    grad = out.means2d.grad
    assert grad is not None and grad.shape == (2, n0, 2)
    g = grad.clone()
    g[..., 0] *= sc.width / 2.0 * 2
    g[..., 1] *= sc.height / 2.0 * 2
    score = g.norm(dim=-1).mean(dim=0)
    thr = float(score.quantile(0.9))                      # densify the top 10 %
    scale_thr = float(torch.exp(P[2].detach()).max(dim=-1).values.median())
    high = score > thr
    small = torch.exp(P[2].detach()).max(dim=-1).values <= scale_thr
    dup_ids, split_ids = torch.nonzero(high & small).reshape(-1), torch.nonzero(high & ~small).reshape(-1)
    before = [p.detach().clone() for p in P]
    ins = InsertUsingImagePlaneGradients(thr, scale_thr)
    n_dup, n_split = ins.step(splats, opt, out, generator=torch.Generator(device=cuda).manual_seed(5))
    assert (n_dup, n_split) == (dup_ids.numel(), split_ids.numel()) and n_dup > 0 and n_split > 0
    assert splats.n == n0 + n_dup + n_split
    new = [p.detach() for p in splats.params]
    for k in range(6):
        assert torch.equal(new[k][:n0], before[k])                       # originals stay
    for k in (0, 1, 2, 3, 4):                                             # duplicates are copies
        assert torch.equal(new[k][n0:n0 + n_dup], before[k][dup_ids])
    assert torch.equal(new[5][n0:], torch.ones(n_dup + n_split, device=cuda))
    for k in (1, 3, 4):                                                   # split: same shape parameters ...
        assert torch.equal(new[k][n0 + n_dup:], before[k][split_ids])
    assert torch.allclose(new[2][n0 + n_dup:], before[2][split_ids] - math.log(1.6), atol=1e-6)
    moved = (new[0][n0 + n_dup:] - before[0][split_ids]).norm(dim=1)
    assert float(moved.max()) > 0 and float(moved.median()) < float(torch.exp(before[2][split_ids]).max()) ** 2 * 10
    assert out.radii.shape == (2, splats.n) and int(out.radii[:, n0:].abs().sum()) == 0
    for p in splats.params:
        st = opt.state[p]
        assert bool((st["exp_avg"][n0:] == 0).all()) and st["exp_avg"].shape == p.shape
