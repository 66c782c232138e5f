"""Densify / prune surgery on the flat splat layout against torch indexing / cat,
which is what the reference does tensor by tensor (gslam/pruning.py:10-53, gslam/insertion.py:27-62)."""
import pytest
import torch

from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats
from gslam_b200.optim import FusedAdam
from gslam_b200.rasterization import rasterization
from gslam_b200.surgery import add_new_splats, prune_using_mask, resize_splats
from gslam_b200.synthetic import make_scene

pytestmark = pytest.mark.gpu


def _setup(cuda, n=5000, capturable=False):
    sc = make_scene(n, 1, "tum", 0.25).to(cuda)
    splats = FlatSplats(sc.params())
    opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(splats.params, PARAM_NAMES)],
                    capturable=capturable)
    g = torch.Generator(device="cpu").manual_seed(3)
    for _ in range(3):                                   # non-trivial Adam moments
        for p in splats.params:
            p.grad = torch.randn(p.shape, generator=g).to(cuda)
        opt.step()
    return sc, splats, opt


def _snapshot(splats, opt):
    return ([p.detach().clone() for p in splats.params],
            [opt.state[p]["exp_avg"].clone() for p in splats.params],
            [opt.state[p]["exp_avg_sq"].clone() for p in splats.params],
            [float(opt.state[p]["step"]) for p in splats.params])


@pytest.mark.parametrize("capturable", [False, True])
def test_prune_then_append_matches_torch_indexing(cuda, capturable):
    sc, splats, opt = _setup(cuda, capturable=capturable)
    P, M, V, S = _snapshot(splats, opt)
    n = splats.n
    keep = torch.rand(n, generator=torch.Generator().manual_seed(1)).to(cuda) > 0.37
    extra = [torch.arange(n, device=cuda)]
    n_pruned = prune_using_mask(splats, opt, keep, per_gaussian_params=extra)
    assert n_pruned == int((~keep).sum()) and splats.n == n - n_pruned
    assert torch.equal(extra[0], torch.arange(n, device=cuda)[keep])
    for k, p in enumerate(splats.params):
        assert p.requires_grad and p.is_leaf
        assert p._base is splats.flat or p.data_ptr() >= splats.flat.data_ptr()
        assert torch.equal(p.detach(), P[k][keep])
        st = opt.state[p]
        assert torch.equal(st["exp_avg"], M[k][keep]) and torch.equal(st["exp_avg_sq"], V[k][keep])
        assert float(st["step"]) == S[k]
        assert opt.param_groups[k]["params"][0] is p and opt.param_groups[k]["lr"] == DEFAULT_LRS[PARAM_NAMES[k]]
    assert len(opt.state) == 6

    # append (gslam/insertion.py:27-62): rows are concatenated, new moments are zero
    add = make_scene(321, 1, "tum", 0.25, seed=9).to(cuda)
    new = dict(zip(PARAM_NAMES, add.params()))
    P2, M2, V2, _ = _snapshot(splats, opt)
    assert add_new_splats(splats, opt, new) == 321
    for k, p in enumerate(splats.params):
        assert torch.equal(p.detach(), torch.cat([P2[k], new[PARAM_NAMES[k]]]))
        st = opt.state[p]
        assert torch.equal(st["exp_avg"][:-321], M2[k]) and bool((st["exp_avg"][-321:] == 0).all())
        assert torch.equal(st["exp_avg_sq"][:-321], V2[k]) and bool((st["exp_avg_sq"][-321:] == 0).all())
    # the flat layout survived: the parameters are the leading rows of six blocks of ONE buffer, one render +
    # backward fills a single [15 N] gradient buffer, and Adam steps
    assert all(p.data_ptr() == b.data_ptr() for p, b in zip(splats.params, splats.blocks()))
    out = rasterization(*splats.params[:5], sc.viewmats, sc.Ks, sc.width, sc.height, packed=False,
                        render_mode="RGB+D", log_uncertainties=splats.params[5],
                        backgrounds=torch.zeros(1, 3, device=cuda))
    out.render.sum().backward()
    g0 = splats.params[0].grad
    assert all(g.untyped_storage().data_ptr() == g0.untyped_storage().data_ptr() for g in (p.grad for p in splats.params))
    before = splats.flat.clone()
    opt.step()
    assert not torch.equal(before, splats.flat)


def test_prune_and_append_in_one_launch_and_edge_masks(cuda):
    sc, splats, opt = _setup(cuda, n=1000)
    P, M, V, _ = _snapshot(splats, opt)
    keep = torch.zeros(1000, dtype=torch.bool, device=cuda)
    keep[::7] = True
    add = make_scene(50, 1, "tum", 0.25, seed=4).to(cuda)
    new = dict(zip(PARAM_NAMES, add.params()))
    assert resize_splats(splats, opt, keep_mask=keep, new_params=new) == int(keep.sum()) + 50
    for k, p in enumerate(splats.params):
        assert torch.equal(p.detach(), torch.cat([P[k][keep], new[PARAM_NAMES[k]]]))
        assert torch.equal(opt.state[p]["exp_avg"][:-50], M[k][keep])
    # keep-nothing mask is a no-op (gslam/pruning.py:16-17); keep-everything prunes zero
    n = splats.n
    assert prune_using_mask(splats, opt, torch.zeros(n, dtype=torch.bool, device=cuda)) == 0 and splats.n == n
    assert prune_using_mask(splats, opt, torch.ones(n, dtype=torch.bool, device=cuda)) == 0 and splats.n == n
    # without optimizer state (fresh optimizer): only the parameters move
    sc2 = make_scene(500, 1, "tum", 0.25).to(cuda)
    s2 = FlatSplats(sc2.params())
    keep2 = torch.rand(500, device=cuda) > 0.5
    ref = [p.detach()[keep2] for p in s2.params]
    prune_using_mask(s2, None, keep2)
    assert all(torch.equal(a.detach(), b) for a, b in zip(s2.params, ref))


def test_prune_round_criteria(cuda):
    """gslam/pruning.py:61-139 + gslam/backend.py:364-391 on the flat layout."""
    from gslam_b200.pruning import (ill_conditioned_mask, large_footprint_mask, low_opacity_mask,
                                    prune_round, young_and_unseen_mask)
    sc, splats, opt = _setup(cuda, n=4000)
    n = splats.n
    g = torch.Generator().manual_seed(2)
    radii = torch.randint(0, 60, (3, n), generator=g).to(torch.int32).to(cuda)
    n_touched = (torch.randint(0, 5, (3, n), generator=g) * (torch.rand(3, n, generator=g) > 0.3)).to(cuda)
    lo = low_opacity_mask(splats.params[3].detach(), 0.3)
    big = large_footprint_mask(radii, 55)
    ill = ill_conditioned_mask(radii, n_touched, 1)
    assert torch.equal(big, radii.max(0).values > 55)
    assert torch.equal(ill, ((radii > 0) & (n_touched == 0)).sum(0) > 1)
    ages = torch.randint(0, 10, (n,), generator=g).to(cuda)
    vis = torch.randint(0, 4, (n,), generator=g).to(cuda)
    assert torch.equal(young_and_unseen_mask(ages, vis, 9, 2), (ages > 6) & (vis < 2))
    expect_remove = lo | big | ill
    kept_means = splats.params[0].detach()[~expect_remove].clone()
    removed = prune_round(splats, opt, radii, n_touched, 0.3, 55, max_unproductive_views=1)
    assert removed == int(expect_remove.sum()) and 0 < removed < n
    assert torch.equal(splats.params[0].detach(), kept_means)
