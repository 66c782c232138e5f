"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle on identical seeded inputs.

Tolerances (BASELINE.json north_star): tile counts / sorted keys / tile ranges
bit-exact on identical projected inputs; images 1e-4 relative; gradients 1e-3
relative (with small absolute floors for values near zero).
"""
import math

import pytest
import torch

from gslam_b200 import ops
from gslam_b200.rasterization import get_new_splat_depth, rasterization
from gslam_b200.synthetic import make_scene
from oracle import torch_ref as R

pytestmark = pytest.mark.gpu

IMG_RTOL, IMG_ATOL = 1e-4, 1e-5
GRAD_RTOL = 1e-3


def rel_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-12))


def oracle_projection(sc):
    scales, opac, colors, betas = R.activations(sc.log_scales, sc.logit_opacities, sc.logit_colors,
                                                sc.log_uncertainties)
    radii, m2, d, con = R.fully_fused_projection(sc.means, sc.quats, scales, sc.viewmats, sc.Ks,
                                                 sc.width, sc.height)
    return radii, m2, d, con, opac, colors, betas


def gpu_project(sc, dev):
    g = sc.to(dev)
    return ops.project(g.means, g.quats, g.log_scales, g.logit_opacities, g.logit_colors,
                       g.log_uncertainties, g.viewmats, g.Ks, g.width, g.height)


@pytest.mark.parametrize("n,c,scale", [(3000, 1, 0.5), (2000, 3, 0.25), (1, 1, 0.25), (257, 2, 0.25)])
def test_projection_forward(cuda, n, c, scale):
    sc = make_scene(n, c, "tum", scale)
    radii, m2, d, con, opac, colors, betas = oracle_projection(sc)
    g_radii, g_m2, g_rec, g_tiles, g_counts = gpu_project(sc, cuda)
    g_radii, g_m2, g_rec, g_tiles = g_radii.cpu(), g_m2.cpu(), g_rec.cpu(), g_tiles.cpu()
    # radii may differ by one where 3*sqrt(lambda) lands within an ulp of an integer
    same = g_radii == radii
    assert same.float().mean() > 0.999
    v = same & (radii > 0)
    torch.testing.assert_close(g_m2[v], m2[v], rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(g_rec[..., 0:2][v], m2[v], rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(g_rec[..., 9][v], d[v], rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(g_rec[..., 3:6][v], con[v], rtol=2e-4, atol=1e-6)
    torch.testing.assert_close(g_rec[..., 2][v], opac.expand_as(radii)[v], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(g_rec[..., 6:9][v], colors.expand(c, -1, -1)[v], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(g_rec[..., 10][v], betas.expand_as(radii)[v], rtol=1e-5, atol=1e-6)
    assert (g_rec[~(g_radii > 0)] == 0).all()
    # tile counts are bit-exact given the GPU's own means2d / radii
    th, tw = math.ceil(sc.height / 16), math.ceil(sc.width / 16)
    tiles_ref, _, _ = R.isect_tiles(g_m2, g_radii, g_rec[..., 9], 16, tw, th, sort=False)
    assert torch.equal(g_tiles, tiles_ref)
    assert int(g_counts[..., 0].sum()) == int(tiles_ref.sum())


# the scatter picks its chunk by launch size: 32 (4000 x 1, 1500 x 4, 5 x 9), 64 (9000 x 1),
# 256 (20000 x 1, 3000 x 8), 2048 (140000 x 1)
@pytest.mark.parametrize("n,c,scale", [(4000, 1, 0.5), (1500, 4, 0.25), (0, 1, 0.25), (5, 9, 0.25),
                                       (9000, 1, 0.25), (20000, 1, 0.25), (3000, 8, 0.2), (140000, 1, 0.1)])
def test_isect_bit_exact(cuda, n, c, scale):
    """Counts, sorted keys, values and tile ranges are bit-exact on identical projected inputs."""
    sc = make_scene(max(n, 1), c, "tum", scale)
    g_radii, g_m2, g_rec, g_tiles, g_counts = gpu_project(sc, cuda)
    if n == 0:
        g_radii.zero_(); g_tiles.zero_(); g_counts.zero_()
    th, tw = math.ceil(sc.height / 16), math.ceil(sc.width / 16)
    ids, fids = ops.isect_tiles(g_m2, g_radii, g_rec, g_tiles, tw, th)
    offs = ops.isect_offset_encode(ids, c, tw, th)
    tiles_ref, ids_ref, fids_ref = R.isect_tiles(g_m2.cpu(), g_radii.cpu(), g_rec[..., 9].cpu(), 16, tw, th)
    offs_ref = R.isect_offset_encode(ids_ref, c, tw, th)
    assert torch.equal(g_tiles.cpu(), tiles_ref)
    assert torch.equal(ids.cpu(), ids_ref)
    assert torch.equal(fids.cpu(), fids_ref)
    assert torch.equal(offs.cpu(), offs_ref)
    # unsorted emission order too
    ids_u, fids_u = ops.isect_tiles(g_m2, g_radii, g_rec, g_tiles, tw, th, sort=False)
    _, ids_ur, fids_ur = R.isect_tiles(g_m2.cpu(), g_radii.cpu(), g_rec[..., 9].cpu(), 16, tw, th, sort=False)
    assert torch.equal(ids_u.cpu(), ids_ur) and torch.equal(fids_u.cpu(), fids_ur)
    # the tile-binned product path (counting sort + per-tile shared-memory sort) gives the same bits
    ids_b, fids_b, offs_b = ops.bin_tiles(g_m2, g_radii, g_rec, g_counts, tw, th)
    assert torch.equal(ids_b.cpu(), ids_ref) and torch.equal(fids_b.cpu(), fids_ref)
    assert torch.equal(offs_b[:-1].cpu().view(c, th, tw), offs_ref)
    assert int(offs_b[-1]) == ids_ref.numel()
    assert int(ops.last_overflow.item()) == 0


def test_bin_tiles_long_segments_and_capacity(cuda):
    """Segments longer than the 8192-entry shared-memory sort, and the no-sync capacity mode."""
    sc = make_scene(60000, 1, "tum", 0.1, log_scale_mean=math.log(0.05))   # 64x48 px: 12 tiles, huge splats
    g_radii, g_m2, g_rec, g_tiles, g_counts = gpu_project(sc, cuda)
    th, tw = math.ceil(sc.height / 16), math.ceil(sc.width / 16)
    _, ids_ref, fids_ref = R.isect_tiles(g_m2.cpu(), g_radii.cpu(), g_rec[..., 9].cpu(), 16, tw, th)
    assert int(g_counts.max()) > 2048
    ids_b, fids_b, offs_b = ops.bin_tiles(g_m2, g_radii, g_rec, g_counts, tw, th)
    assert torch.equal(ids_b.cpu(), ids_ref) and torch.equal(fids_b.cpu(), fids_ref)
    I = ids_ref.numel()
    try:
        ops.isect_capacity = I + 1000
        ids_c, fids_c, offs_c = ops.bin_tiles(g_m2, g_radii, g_rec, g_counts, tw, th)
        assert ids_c.numel() == I + 1000 and int(ops.last_overflow.item()) == 0
        assert int(ops.last_n_isects.item()) == I
        assert torch.equal(ids_c[:I].cpu(), ids_ref) and torch.equal(fids_c[:I].cpu(), fids_ref)
        ops.isect_capacity = I // 2
        ops.bin_tiles(g_m2, g_radii, g_rec, g_counts, tw, th)
        assert int(ops.last_overflow.item()) == 1
    finally:
        ops.isect_capacity = None


def test_overflowed_capacity_step_stays_in_bounds(cuda):
    """A step with more intersections than its (graph-captured) key buffers hold raises the overflow
    flag and renders a truncated image, but forward and backward never read past the buffers."""
    sc = make_scene(30000, 2, "tum", 0.5).to(cuda)
    ps = [p.clone().requires_grad_(True) for p in sc.params()]
    vm = sc.viewmats.clone().requires_grad_(True)
    bg = torch.zeros(2, 3, device=cuda)

    def run():
        out = rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, sc.Ks, sc.width, sc.height, packed=False,
                            backgrounds=bg, render_mode="RGB+D", log_uncertainties=ps[5])
        (out.render.sum() + out.alphas.sum()).backward()
        return out
    full = run()
    I = int(ops.last_n_isects.item())
    try:
        for cap in (I // 3, 64, I - 1):
            ops.isect_capacity = cap
            out = run()
            torch.cuda.synchronize()
            assert int(ops.last_overflow.item()) == 1 and out.flatten_ids.numel() == cap
            assert torch.isfinite(out.render).all() and torch.isfinite(vm.grad).all()
        ops.isect_capacity = I + 7
        out = run()
        assert int(ops.last_overflow.item()) == 0
        assert torch.allclose(out.render, full.render, atol=1e-6)
    finally:
        ops.isect_capacity = None


def test_sort_random_keys_stable(cuda):
    g = torch.Generator().manual_seed(3)
    for n, bits in [(1, 40), (1000, 44), (100_000, 49), (300_001, 33)]:
        keys = torch.randint(0, 1 << 20, (n,), generator=g, dtype=torch.int64) << (bits - 20)
        keys |= torch.randint(0, 4, (n,), generator=g, dtype=torch.int64)   # many duplicates
        vals = torch.arange(n, dtype=torch.int32)
        ko, vo = ops.sort_pairs(keys.to(cuda), vals.to(cuda), bits)
        kr, order = torch.sort(keys, stable=True)
        assert torch.equal(ko.cpu(), kr)
        assert torch.equal(vo.cpu(), vals[order])


def _composite_inputs(sc, dev):
    """Oracle projection + binning, moved to the GPU as a splat record."""
    radii, m2, d, con, opac, colors, betas = oracle_projection(sc)
    C, N = radii.shape
    th, tw = math.ceil(sc.height / 16), math.ceil(sc.width / 16)
    _, ids, fids = R.isect_tiles(m2, radii, d, 16, tw, th)
    offs = R.isect_offset_encode(ids, C, tw, th)
    colors5 = torch.cat([colors.expand(C, -1, -1), d[..., None], betas.expand(C, -1)[..., None]], -1)
    opac2 = opac.expand(C, -1).contiguous()
    rec = torch.zeros(C, N, 12)
    rec[..., 0:2] = m2; rec[..., 2] = opac2; rec[..., 3:6] = con; rec[..., 6:11] = colors5
    return dict(m2=m2, con=con, colors5=colors5, opac=opac2, fids=fids, offs=offs, rec=rec,
                radii=radii, tw=tw, th=th)


@pytest.mark.parametrize("n,c,scale,bg", [(3000, 1, 0.5, True), (1200, 2, 0.25, False),
                                          (20000, 1, 0.25, True)])
def test_composite_forward_and_backward(cuda, n, c, scale, bg):
    """K6/K7 on the oracle's own sorted lists: images, alphas, n_touched, last ids, all gradients."""
    sc = make_scene(n, c, "tum", scale)
    I = _composite_inputs(sc, cuda)
    bgs = torch.tensor([[0.1, 0.2, 0.3, 0.0, math.e]]).repeat(c, 1) if bg else None
    m2 = I["m2"].clone().requires_grad_(True)
    con = I["con"].clone().requires_grad_(True)
    col = I["colors5"].clone().requires_grad_(True)
    op = I["opac"].clone().requires_grad_(True)
    ref, ref_a, ref_nt, st = R.rasterize_to_pixels(m2, con, col, op, sc.width, sc.height, 16, I["offs"],
                                                   I["fids"], backgrounds=bgs, return_stats=True)
    g = torch.Generator().manual_seed(5)
    w_img = torch.rand(ref.shape, generator=g)
    w_a = torch.rand(ref_a.shape, generator=g)
    ((ref * w_img).sum() + (ref_a * w_a).sum()).backward()

    rec = I["rec"].to(cuda).requires_grad_(True)
    m2g = I["m2"].to(cuda).requires_grad_(True)
    out, out_a, nt = ops.composite(rec, m2g, I["fids"].to(cuda), I["offs"].to(cuda),
                                   bgs.to(cuda) if bg else None, sc.width, sc.height)
    torch.testing.assert_close(out.cpu(), ref.detach(), rtol=IMG_RTOL, atol=IMG_ATOL)
    torch.testing.assert_close(out_a.cpu(), ref_a.detach(), rtol=IMG_RTOL, atol=IMG_ATOL)
    assert rel_err(out, ref.detach()) < 1e-5
    # n_touched: integer counts; pairs exactly at a threshold may flip, allow a 1e-3 fraction
    diff = (nt.cpu() - ref_nt).abs().sum().item()
    assert diff <= max(2, 1e-3 * ref_nt.sum().item()), diff
    ((out * w_img.to(cuda)).sum() + (out_a * w_a.to(cuda)).sum()).backward()
    v_rec = rec.grad.cpu()
    assert rel_err(m2g.grad, m2.grad) < GRAD_RTOL
    assert rel_err(v_rec[..., 2], op.grad) < GRAD_RTOL
    assert rel_err(v_rec[..., 3:6], con.grad) < GRAD_RTOL
    assert rel_err(v_rec[..., 6:11], col.grad) < GRAD_RTOL


def _full(sc, dev, render_mode="RGB+D", with_unc=True, bg=True):
    """Run oracle and CUDA rasterization() on the same scene with grads to everything."""
    def leafs(s, device):
        ps = [p.clone().to(device).requires_grad_(True) for p in s.params()]
        vm = s.viewmats.clone().to(device).requires_grad_(True)
        return ps, vm
    C = sc.viewmats.shape[0]
    bgs = torch.tensor([[0.05, 0.1, 0.15]]).repeat(C, 1) if bg else None
    ps, vm = leafs(sc, "cpu")
    ref = R.rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, sc.Ks, sc.width, sc.height,
                          backgrounds=bgs, render_mode=render_mode,
                          log_uncertainties=ps[5] if with_unc else None)
    gps, gvm = leafs(sc, dev)
    out = rasterization(gps[0], gps[1], gps[2], gps[3], gps[4], gvm, sc.Ks.to(dev), sc.width, sc.height,
                        packed=False, backgrounds=bgs.to(dev) if bg else None, render_mode=render_mode,
                        log_uncertainties=gps[5] if with_unc else None)
    return ref, (ps, vm), out, (gps, gvm)


@pytest.mark.parametrize("n,c,scale", [(4000, 1, 0.5), (1500, 3, 0.25)])
def test_rasterization_end_to_end(cuda, n, c, scale):
    sc = make_scene(n, c, "tum", scale)
    ref, (ps, vm), out, (gps, gvm) = _full(sc, cuda)
    assert out.rgbs.shape == ref.rgbs.shape and out.alphas.shape == ref.alphas.shape
    assert out.depthmaps.shape == ref.depthmaps.shape and out.betas.shape == ref.betas.shape
    # images: 1e-4 relative (norm-wise) + elementwise with a small floor for radius-rounding pixels
    for a, b in [(out.rgbs, ref.rgbs), (out.alphas, ref.alphas), (out.depthmaps, ref.depthmaps),
                 (out.betas, ref.betas)]:
        assert rel_err(a, b.detach()) < IMG_RTOL
    assert out.n_touched.dtype == torch.int64 and out.radii.dtype == torch.int32
    assert out.isect_offsets.shape == (c, out.tile_height, out.tile_width)
    assert abs(out.isect_ids.numel() - ref.isect_ids.numel()) <= max(4, 1e-3 * ref.isect_ids.numel())
    # backward through a mapping-style loss
    g = torch.Generator().manual_seed(7)
    w = torch.rand(ref.rgbs.shape, generator=g)
    def loss(o, w_):
        return ((o.rgbs * w_).sum() + 0.1 * o.depthmaps.sum() + 0.3 * (o.betas * o.betas).sum()
                + 0.2 * o.alphas.sum())
    loss(ref, w).backward()
    out.means2d.retain_grad()
    loss(out, w.to(cuda)).backward()
    names = ["means", "quats", "log_scales", "logit_opacities", "logit_colors", "log_uncertainties"]
    for name, p, gp in zip(names, ps, gps):
        assert gp.grad is not None, name
        assert rel_err(gp.grad, p.grad) < GRAD_RTOL, (name, rel_err(gp.grad, p.grad))
    assert rel_err(gvm.grad[:, :3], vm.grad[:, :3]) < GRAD_RTOL
    assert out.means2d.grad is not None and out.means2d.grad.shape == (c, n, 2)


def test_tracking_only_camera_gradient(cuda):
    """Splats without grad (gslam/backend.py:515 no_grad_clone): only the viewmat gets a gradient."""
    sc = make_scene(3000, 1, "tum", 0.5)
    ps = [p.clone().requires_grad_(True) for p in sc.params()]
    vm = sc.viewmats.clone().requires_grad_(True)
    ref = R.rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, sc.Ks, sc.width, sc.height,
                          render_mode="RGB+D", log_uncertainties=ps[5], backgrounds=torch.zeros(1, 3))
    (ref.rgbs.square().sum() + ref.betas.sum()).backward()
    g = sc.to(cuda)
    gvm = g.viewmats.clone().requires_grad_(True)
    out = rasterization(g.means, g.quats, g.log_scales, g.logit_opacities, g.logit_colors, gvm, g.Ks,
                        g.width, g.height, packed=False, render_mode="RGB+D",
                        log_uncertainties=g.log_uncertainties, backgrounds=torch.zeros(1, 3, device=cuda))
    (out.rgbs.square().sum() + out.betas.sum()).backward()
    assert rel_err(gvm.grad[:, :3], vm.grad[:, :3]) < GRAD_RTOL
    assert (gvm.grad[:, 3] == 0).all()


def test_render_modes_and_rejections(cuda):
    sc = make_scene(500, 1, "tum", 0.25).to(cuda)
    a = (sc.means, sc.quats, sc.log_scales, sc.logit_opacities, sc.logit_colors, sc.viewmats, sc.Ks,
         sc.width, sc.height)
    o = rasterization(*a, packed=False)   # RGB, no uncertainties, no background
    assert o.depthmaps is None and o.betas is None and o.rgbs.shape[-1] == 3
    o = rasterization(*a, packed=False, render_mode="D")
    assert o.rgbs is None and o.depthmaps is not None
    for kw in [dict(packed=True), dict(packed=False, render_mode="ED"), dict(packed=False, tile_size=8),
               dict(packed=False, absgrad=True), dict(packed=False, camera_model="fisheye")]:
        with pytest.raises(NotImplementedError):
            rasterization(*a, **kw)
    with pytest.raises(RuntimeError):
        c = make_scene(10, 1, "tum", 0.25)
        rasterization(c.means, c.quats, c.log_scales, c.logit_opacities, c.logit_colors, c.viewmats,
                      c.Ks, c.width, c.height, packed=False)


def test_culling_and_empty(cuda):
    """Behind the camera / off screen -> radii 0; empty scene renders the background."""
    sc = make_scene(64, 1, "tum", 0.25)
    sc.means[:32, 2] = -1.0          # behind the near plane
    sc.means[32:, 0] = 1e4           # far off screen
    g = sc.to(cuda)
    bg = torch.tensor([[0.2, 0.4, 0.6]], device=cuda)
    o = rasterization(g.means, g.quats, g.log_scales, g.logit_opacities, g.logit_colors, g.viewmats,
                      g.Ks, g.width, g.height, packed=False, backgrounds=bg, render_mode="RGB+D",
                      log_uncertainties=g.log_uncertainties)
    assert (o.radii == 0).all() and o.isect_ids.numel() == 0 and (o.isect_offsets == 0).all()
    assert (o.alphas == 0).all() and (o.depthmaps == 0).all()
    torch.testing.assert_close(o.rgbs, bg.view(1, 1, 1, 3).expand_as(o.rgbs))
    torch.testing.assert_close(o.betas, torch.full_like(o.betas, math.e))
    # ... and its backward is all zeros (a tracker whose line search looks away from the map gets here)
    ps = [p.clone().requires_grad_(True) for p in g.params()]
    vm = g.viewmats.clone().requires_grad_(True)
    o = rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, g.Ks, g.width, g.height, packed=False,
                      backgrounds=bg, render_mode="RGB+D", log_uncertainties=ps[5])
    (o.render.square().sum() + o.alphas.sum()).backward()
    assert all(p.grad is None or float(p.grad.abs().max()) == 0.0 for p in ps)
    assert vm.grad is None or float(vm.grad.abs().max()) == 0.0


def test_packed_projection(cuda):
    sc = make_scene(800, 3, "tum", 0.25)
    radii, m2, d, *_ = oracle_projection(sc)
    g = sc.to(cuda)
    cam, gid, r, m, dep = get_new_splat_depth(
        {"means": g.means, "quats": g.quats, "scales": g.log_scales}, g.viewmats, g.Ks, g.width, g.height)
    cref, gref = torch.nonzero(radii > 0, as_tuple=True)
    assert cam.dtype == torch.int64 and r.dtype == torch.int32
    # compare on the intersection (radius rounding may flip a borderline cull)
    key = (cam * 800 + gid).cpu(); kref = cref * 800 + gref
    assert len(set(key.tolist()) ^ set(kref.tolist())) <= 2
    common = torch.isin(key, kref)
    sel = torch.isin(kref, key)
    torch.testing.assert_close(m.cpu()[common], m2[cref, gref][sel], rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(dep.cpu()[common], d[cref, gref][sel], rtol=1e-5, atol=1e-5)
    r2, m2u, d2u = get_new_splat_depth(
        {"means": g.means, "quats": g.quats, "scales": g.log_scales}, g.viewmats, g.Ks, g.width,
        g.height, packed=False)
    assert r2.shape == (3, 800) and m2u.shape == (3, 800, 2) and d2u.shape == (3, 800)


@pytest.mark.parametrize("n", [1, 2, 3, 1001, 2503])
def test_backward_with_gaussian_counts_not_multiple_of_four(cuda, n):
    """The flat [15 N] gradient puts the quaternion rows 12 N bytes in: not 16-byte aligned unless
    N % 4 == 0.  Forward and backward must not care."""
    sc = make_scene(n, 2, "tum", 0.25)
    ref, (ps, vm), out, (gps, gvm) = _full(sc, cuda)
    w = torch.rand(ref.rgbs.shape, generator=torch.Generator().manual_seed(3))
    ((ref.rgbs * w).sum() + ref.alphas.sum()).backward()
    ((out.rgbs * w.to(cuda)).sum() + out.alphas.sum()).backward()
    for p, gp in zip(ps[:5], gps[:5]):
        assert gp.grad is not None and bool(torch.isfinite(gp.grad).all())
        if float(p.grad.norm()) > 0:
            assert rel_err(gp.grad, p.grad) < GRAD_RTOL
