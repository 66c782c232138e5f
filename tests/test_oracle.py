"""Known-answer and property tests that pin the CPU oracle (oracle/torch_ref.py).

The reference holds no tests or golden vectors for this path, so the
oracle is pinned by closed forms, by finite differences in float64, by an independent
sequential per-pixel loop, and by torch's own Adam.  CPU only.
"""
import math

import pytest
import torch

from gslam_b200.synthetic import make_scene
from oracle import torch_ref as R


def _one_gaussian(z=2.0, s=0.05, W=64, H=48, f=50.0):
    K = torch.tensor([[[f, 0, W / 2], [0, f, H / 2], [0, 0, 1.0]]])
    means = torch.tensor([[0.0, 0.0, z]])
    quats = torch.tensor([[1.0, 0, 0, 0]])
    scales = torch.full((1, 3), s)
    return means, quats, scales, torch.eye(4)[None], K, W, H


def test_projection_closed_form_isotropic_on_axis():
    means, quats, scales, vm, K, W, H = _one_gaussian()
    radii, m2, d, con = R.fully_fused_projection(means, quats, scales, vm, K, W, H)
    var = (50.0 * 0.05 / 2.0) ** 2 + 0.3                      # (f s / z)^2 + eps2d
    assert torch.allclose(m2[0, 0], torch.tensor([32.0, 24.0]))
    assert abs(d[0, 0].item() - 2.0) < 1e-7
    assert torch.allclose(con[0, 0], torch.tensor([1 / var, 0.0, 1 / var]), atol=1e-6)
    lam = var + math.sqrt(max(0.01, 0.0))                      # b + sqrt(max(0.01, b^2 - det)), b^2 == det
    assert radii[0, 0].item() == math.ceil(3 * math.sqrt(lam))


@pytest.mark.parametrize("mean,expect_visible", [
    ([0.0, 0.0, 0.005], False),     # in front of the near plane (0.01)
    ([0.0, 0.0, -1.0], False),      # behind the camera
    ([100.0, 0.0, 2.0], False),     # far off screen
    ([0.0, 0.0, 2.0], True),
])
def test_projection_culls(mean, expect_visible):
    _, quats, scales, vm, K, W, H = _one_gaussian()
    radii, m2, d, con = R.fully_fused_projection(torch.tensor([mean]), quats, scales, vm, K, W, H)
    assert (radii[0, 0].item() > 0) == expect_visible
    if not expect_visible:
        assert m2.abs().sum() == 0 and d.abs().sum() == 0 and con.abs().sum() == 0


def test_quat_scale_to_covar_is_rotation_invariant_for_isotropic_scales():
    q = torch.randn(5, 4)
    cov = R.quat_scale_to_covar(q, torch.full((5, 3), 0.3))
    assert torch.allclose(cov, 0.09 * torch.eye(3).expand(5, 3, 3), atol=1e-6)
    Rm = R.quat_to_rotmat(q)
    assert torch.allclose(Rm @ Rm.transpose(-1, -2), torch.eye(3).expand(5, 3, 3), atol=1e-5)
    assert torch.allclose(torch.linalg.det(Rm), torch.ones(5), atol=1e-5)


def test_tile_bbox_edges_and_key_layout():
    # centre exactly on a tile corner, radius exactly one tile: covers [0,2) x [0,2)
    m2 = torch.tensor([[[16.0, 16.0], [8.0, 8.0], [70.0, 40.0]]])
    radii = torch.tensor([[16, 8, 0]], dtype=torch.int32)
    depths = torch.tensor([[1.5, 0.5, 9.0]])
    tpg, ids, fids = R.isect_tiles(m2, radii, depths, 16, 4, 3, sort=False)
    assert tpg.tolist() == [[4, 1, 0]]
    tile_n_bits, cam_n_bits = R.tile_bits(4, 3, 1)
    assert (tile_n_bits, cam_n_bits) == (4, 1)
    tiles = ((ids >> 32) & ((1 << tile_n_bits) - 1)).tolist()
    assert tiles == [0, 1, 4, 5, 0]                                  # rows then columns, c-major / n order
    dbits = (ids & 0xFFFFFFFF).to(torch.int32).view(torch.float32)
    assert dbits.tolist() == [1.5, 1.5, 1.5, 1.5, 0.5]
    assert fids.tolist() == [0, 0, 0, 0, 1]
    # sorted: tile 0 holds the nearer splat (id 1) first
    _, ids_s, fids_s = R.isect_tiles(m2, radii, depths, 16, 4, 3)
    assert fids_s.tolist() == [1, 0, 0, 0, 0]
    assert torch.all(ids_s[1:] >= ids_s[:-1])
    off = R.isect_offset_encode(ids_s, 1, 4, 3)
    assert off.view(-1).tolist() == [0, 2, 3, 3, 3, 4, 5, 5, 5, 5, 5, 5]   # empty tiles take the next start


def test_keys_sort_by_camera_then_tile_then_depth():
    sc = make_scene(300, 3, "tum", 0.1)
    scales, opac, colors, betas = R.activations(sc.log_scales, sc.logit_opacities, sc.logit_colors)
    radii, m2, d, _ = R.fully_fused_projection(sc.means, sc.quats, scales, sc.viewmats, sc.Ks, sc.width, sc.height)
    tw, th = math.ceil(sc.width / 16), math.ceil(sc.height / 16)
    tpg, ids, fids = R.isect_tiles(m2, radii, d, 16, tw, th)
    tb, _ = R.tile_bits(tw, th, 3)
    cam, tile = ids >> (32 + tb), (ids >> 32) & ((1 << tb) - 1)
    assert torch.equal(cam, fids.long() // 300)                       # key camera == value camera
    lin = cam * tw * th + tile
    assert torch.all(lin[1:] >= lin[:-1])
    same = lin[1:] == lin[:-1]
    dd = d.reshape(-1)[fids.long()]
    assert torch.all(dd[1:][same] >= dd[:-1][same])                   # depth ascending inside a tile
    assert int(tpg.sum()) == ids.numel()


def _sequential_composite(m2, con, col, op, W, H, offs, fids, bg, min_T=0.5):
    """The per-pixel loop exactly as the CUDA reference runs it (slow; small inputs only)."""
    C, N = op.shape
    D = col.shape[-1]
    th, tw = offs.shape[1:]
    flat = offs.reshape(-1).tolist() + [fids.numel()]
    out = torch.zeros(C, H, W, D)
    alpha_out = torch.zeros(C, H, W)
    touched = torch.zeros(C * N, dtype=torch.int64)
    m2f, conf, colf, opf = m2.reshape(-1, 2), con.reshape(-1, 3), col.reshape(-1, D), op.reshape(-1)
    for c in range(C):
        for y in range(H):
            for x in range(W):
                t = (c * th + y // 16) * tw + x // 16
                T = 1.0
                acc = torch.zeros(D)
                for k in range(flat[t], flat[t + 1]):
                    g = int(fids[k])
                    dx, dy = float(m2f[g, 0]) - (x + 0.5), float(m2f[g, 1]) - (y + 0.5)
                    a, b, cc = conf[g].tolist()
                    sigma = 0.5 * (a * dx * dx + cc * dy * dy) + b * dx * dy
                    alpha = min(0.999, float(opf[g]) * math.exp(-sigma))
                    if sigma < 0 or alpha < 1 / 255:
                        continue
                    nT = T * (1 - alpha)
                    if nT <= 1e-4:
                        break
                    acc += colf[g] * alpha * T
                    if T > min_T:
                        touched[g] += 1
                    T = nT
                out[c, y, x] = acc + T * bg[c]
                alpha_out[c, y, x] = 1 - T
    return out, alpha_out, touched.reshape(C, N)


def test_vectorised_compositing_equals_sequential_loop():
    sc = make_scene(150, 2, "tum", 0.05, log_scale_mean=math.log(0.04))     # 32x24 px, opaque-ish splats
    sc.logit_opacities += 2.0                                                # force early saturation
    scales, opac, colors, betas = R.activations(sc.log_scales, sc.logit_opacities, sc.logit_colors,
                                                sc.log_uncertainties)
    radii, m2, d, con = R.fully_fused_projection(sc.means, sc.quats, scales, sc.viewmats, sc.Ks, sc.width, sc.height)
    tw, th = math.ceil(sc.width / 16), math.ceil(sc.height / 16)
    _, ids, fids = R.isect_tiles(m2, radii, d, 16, tw, th)
    offs = R.isect_offset_encode(ids, 2, tw, th)
    col = torch.cat([colors.expand(2, -1, -1), d[..., None]], -1)
    op = opac.expand(2, -1)
    bg = torch.tensor([[0.1, 0.2, 0.3, 0.0]] * 2)
    out, a, nt, st = R.rasterize_to_pixels(m2, con, col, op, sc.width, sc.height, 16, offs, fids,
                                           backgrounds=bg, return_stats=True)
    ref, ref_a, ref_nt = _sequential_composite(m2, con, col, op, sc.width, sc.height, offs, fids, bg)
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(a[..., 0], ref_a, rtol=1e-4, atol=1e-5)
    assert torch.equal(nt.long(), ref_nt)
    assert (a.max() > 0.999) and st["P"] < fids.numel() * 256          # some pixels really stopped early


def test_two_overlapping_gaussians_analytic_pixel():
    W, H = 16, 16
    m2 = torch.tensor([[[8.5, 8.5], [8.5, 8.5]]])
    con = torch.tensor([[[0.5, 0.0, 0.5], [0.5, 0.0, 0.5]]])
    col = torch.tensor([[[1.0, 0.0, 0.0], [0.0, 1.0, 0.0]]])
    op = torch.tensor([[0.6, 0.5]])
    offs = torch.zeros(1, 1, 1, dtype=torch.int32)
    out, a, nt = R.rasterize_to_pixels(m2, con, col, op, W, H, 16, offs, torch.tensor([0, 1], dtype=torch.int32))
    px = out[0, 8, 8]                                                 # pixel centre (8.5, 8.5): sigma = 0
    assert torch.allclose(px, torch.tensor([0.6, 0.4 * 0.5, 0.0]), atol=1e-6)
    assert abs(a[0, 8, 8, 0].item() - (1 - 0.4 * 0.5)) < 1e-6
    # the 2nd splat is not counted where the 1st left T <= 0.5 (the few pixels nearest the centre)
    assert nt[0, 0] > nt[0, 1] > 0


def test_gradcheck_float64_all_inputs_including_pose():
    sc = make_scene(12, 2, "tum", 0.05, dtype=torch.float64, log_scale_mean=math.log(0.05))
    gt = torch.rand(2, sc.height, sc.width, 3, dtype=torch.float64)
    Rt = sc.viewmats[1].clone()

    def f(means, quats, ls, lo, lc, lu, dR, dt):
        vm = torch.stack([sc.viewmats[0], R.pose_zhou(Rt, dR, dt)])
        o = R.rasterization(means, quats, ls, lo, lc, vm, sc.Ks, sc.width, sc.height,
                            backgrounds=torch.zeros(2, 3, dtype=torch.float64), render_mode="RGB+D",
                            log_uncertainties=lu)
        return ((o.rgbs - gt).square().sum(-1) / (2 * o.betas.square())).mean() + 0.1 * o.depthmaps.mean() \
            + 0.1 * o.alphas.mean()

    inputs = [p.clone().requires_grad_(True) for p in sc.params()]
    inputs += [torch.zeros(6, dtype=torch.float64, requires_grad=True),
               torch.zeros(3, dtype=torch.float64, requires_grad=True)]
    assert torch.autograd.gradcheck(f, inputs, eps=1e-6, atol=1e-6, rtol=1e-3, nondet_tol=0.0)


def test_pose_zhou_identity_and_orthonormal():
    Rt = torch.eye(4)
    Rt[:3, 3] = torch.tensor([0.1, -0.2, 0.3])
    assert torch.allclose(R.pose_zhou(Rt, torch.zeros(6), torch.zeros(3)), Rt)
    M = R.rotation_6d_to_matrix(torch.randn(4, 6))
    assert torch.allclose(M @ M.transpose(-1, -2), torch.eye(3).expand(4, 3, 3), atol=1e-5)


def test_ssim_properties():
    a = torch.rand(2, 3, 32, 40)
    assert abs(R.fused_ssim(a, a).item() - 1.0) < 1e-6
    assert abs(R.fused_ssim(a, a, padding="valid").item() - 1.0) < 1e-6
    b = torch.rand(2, 3, 32, 40)
    assert abs(R.fused_ssim(a, b).item() - R.fused_ssim(b, a).item()) < 1e-6      # symmetric
    assert R.fused_ssim(a, b).item() < 0.2
    w = R.gaussian_window()
    assert abs(w.sum().item() - 1) < 1e-6 and w.argmax() == 5
    # 'valid' is the 'same' map without its 5-pixel border
    m = R.ssim_map(a, b)
    assert abs(R.fused_ssim(a, b, "valid").item() - m[:, :, 5:-5, 5:-5].mean().item()) < 1e-7


def test_adam_matches_torch_optim():
    p = torch.randn(7, 3)
    ref = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=1.6e-3)
    m, v, x = torch.zeros_like(p), torch.zeros_like(p), p.clone()
    for step in range(1, 6):
        g = torch.randn(7, 3)
        ref.grad = g.clone()
        opt.step()
        x, m, v = R.adam_step(x, g, m, v, step, 1.6e-3)
        torch.testing.assert_close(x, ref.detach(), rtol=1e-6, atol=1e-7)


def test_opacity_decay_mask():
    lo = torch.ones(4)
    radii = torch.tensor([[1, 0, 3, 0], [2, 0, 0, 5], [0, 0, 4, 0]])
    assert R.opacity_decay(lo, radii, 0.5).tolist() == [0.5, 1.0, 0.5, 1.0]       # seen in > 1 view only


def test_quat_scale_to_covar_preci_shim_matches_oracle():
    """The torch shim of gsplat.quat_scale_to_covar_preci (gslam/insertion.py:88) vs the oracle's covariance."""
    from gslam_b200.rasterization import quat_scale_to_covar_preci
    g = torch.Generator().manual_seed(3)
    q = torch.rand(64, 4, generator=g)
    s = torch.rand(64, 3, generator=g) + 0.05
    covar, preci = quat_scale_to_covar_preci(q, s)
    assert torch.allclose(covar, R.quat_scale_to_covar(q, s), atol=1e-6)
    assert torch.allclose(covar @ preci, torch.eye(3).expand(64, 3, 3), atol=1e-4)


def test_sh_basis_is_orthonormal_and_degree0_constant():
    """SH extension of the oracle: the 16 basis functions are orthonormal on the sphere
    (Monte-Carlo quadrature), and a degree-0 colour does not depend on the view."""
    g = torch.Generator().manual_seed(0)
    d = torch.nn.functional.normalize(torch.randn(400_000, 3, generator=g, dtype=torch.float64), dim=-1)
    Y = R.sh_basis(3, d)
    gram = (Y.T @ Y) * (4 * math.pi / d.shape[0])
    assert torch.allclose(gram, torch.eye(16, dtype=torch.float64), atol=2e-2)
    means = torch.randn(50, 3, generator=g)
    vm = torch.eye(4).repeat(3, 1, 1)
    vm[1, :3, 3] = torch.tensor([0.3, -0.2, 0.5])
    vm[2, :3, 3] = torch.tensor([-1.0, 0.4, 0.1])
    c = torch.randn(50, 16, 3, generator=g)
    col0 = R.sh_colors(0, c, means, vm)
    assert torch.allclose(col0[0], col0[1]) and torch.allclose(col0[0], col0[2])
    col3 = R.sh_colors(3, c, means, vm)
    assert not torch.allclose(col3[0], col3[2])
    assert float(col3.min()) >= 0.0
