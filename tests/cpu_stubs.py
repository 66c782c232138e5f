"""Oracle-based stand-ins for the CUDA pieces of `gslam_b200.mapping.MappingStep`, for tests of its HOST
logic on the CPU (gloo): the renderer and the image loss come from oracle/torch_ref.py, the gradient rows
are packed from ordinary autograd gradients, and the fused update (isotropic regulariser -> Adam -> opacity
decay) is restated with torch ops.  Test infrastructure only: the product package has no CPU branch.
"""
import torch

from oracle import torch_ref as R


def oracle_render(*a, **kw):
    out = R.rasterization(*a, **kw)
    out.render = torch.cat([out.rgbs, out.depthmaps[..., None], out.betas[..., None]], dim=-1)
    return out


def oracle_loss(render, gt, exposure, w, mode):
    photo = R.mapping_photometric_loss(R.exposure_affine(render[..., :3], exposure), gt, render[..., 4])
    ssim = 1.0 - R.fused_ssim(render[..., :3].permute(0, 3, 1, 2), gt.permute(0, 3, 1, 2), padding="valid")
    return (1 - w) * photo + w * ssim, photo.detach()


@torch.no_grad()
def collect_grads(params, radii, rows, lay):
    """What gsb_project_bwd's row layout does: p.grad of the six parameters and the visibility count of
    this rank's views into rows[n, :]."""
    n = params[0].shape[0]
    for p, c, w in zip(params, lay.cols, lay.widths):
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        rows[:n, c:c + w] = g.reshape(n, w)
    rows[:n, lay.vis] = (radii > 0).sum(dim=0).to(rows.dtype)


def isotropic_grad(log_scales, visible, weight):
    """(gradient [N,3], unweighted loss) of weight * sum_visible |exp(s) - exp(mean_axis s).detach()|."""
    s = log_scales.detach()
    e = s.exp()
    m = s.mean(dim=1, keepdim=True).exp()
    vis = visible[:, None].to(s.dtype)
    return weight * torch.sign(e - m) * e * vis, ((e - m).abs() * vis).sum()


class TorchRowsAdam:
    """torch.optim.Adam semantics on row ranges: one step counter, per-parameter learning rates."""

    def __init__(self, params, lrs, betas=(0.9, 0.999), eps=1e-8):
        self.lrs, self.betas, self.eps = list(lrs), betas, eps
        self.m = [torch.zeros_like(p) for p in params]
        self.v = [torch.zeros_like(p) for p in params]
        self.t = 0

    @torch.no_grad()
    def update(self, step, rows, lay, n0, n1, vis_out):
        """The `update_fn` of MappingStep: gsb_map_update restated."""
        self.t += 1
        b1, b2 = self.betas
        bc1, bc2 = 1 - b1 ** self.t, 1 - b2 ** self.t
        P = step.splats.params
        vis = rows[n0:n1, lay.vis]
        vis_out[n0:n1] = vis.to(vis_out.dtype)
        iso = torch.zeros(1)
        for k, (p, c, w) in enumerate(zip(P, lay.cols, lay.widths)):
            g = rows[n0:n1, c:c + w].reshape(p[n0:n1].shape).clone()
            if k == 2:
                gi, li = isotropic_grad(p[n0:n1], vis > 0, step.iso_w)
                g += gi
                iso = li.view(1)
            m, v = self.m[k][n0:n1], self.v[k][n0:n1]
            m.mul_(b1).add_(g, alpha=1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            upd = (self.lrs[k] / bc1) * m / (v.sqrt() / bc2 ** 0.5 + self.eps)
            p[n0:n1] -= upd
            if k == 3:
                p[n0:n1] *= torch.where(vis > 1, step.decay, 1.0)
        return iso
