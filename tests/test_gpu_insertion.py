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
