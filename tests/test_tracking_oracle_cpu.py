"""The tracking algorithm end to end on the CPU oracle (no GPU): pose chain (gslam/primitives.py:
15-36, 82-92) -> oracle rasterization -> exposure + active-NeRF tracking loss (gslam/frontend.py:
113-129, 632-636) as the objective of gslam_b200.lbfgs, the optimizer the CUDA tracker drives with
graph replays.  Pins the host-side half of gslam_b200.tracking independently of the kernels: the
L-BFGS restatement takes the same evaluations to the same pose and loss as torch.optim.LBFGS on
this (deterministic, float32) objective."""
import numpy as np
import torch

from gslam_b200.lbfgs import lbfgs_minimize
from gslam_b200.synthetic import make_scene, smooth_trajectory
from oracle import torch_ref as R


def _objective(sc, gt, Rt):
    ps = [p.detach() for p in sc.params()]
    bg = torch.zeros(1, 3)

    def f(theta: torch.Tensor) -> torch.Tensor:
        vm = R.pose_zhou(Rt, theta[0:6], theta[6:9]).view(1, 4, 4)
        out = R.rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, sc.Ks, sc.width, sc.height,
                              backgrounds=bg, render_mode="RGB+D", log_uncertainties=ps[5])
        rgb = R.exposure_affine(out.rgbs, theta[9:11].view(1, 2))
        return R.tracking_loss(gt, rgb[0], out.betas[0])
    return f


def test_tracking_schedule_on_the_oracle_objective_matches_torch_lbfgs():
    torch.set_num_threads(4)
    sc = make_scene(400, 1, "tum", 0.125)                      # 80 x 60
    true = smooth_trajectory(13)[12]
    ps = [p.detach() for p in sc.params()]
    with torch.no_grad():
        gt = R.rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], true.view(1, 4, 4), sc.Ks, sc.width, sc.height,
                             backgrounds=torch.zeros(1, 3), render_mode="RGB+D",
                             log_uncertainties=ps[5]).rgbs[0].clamp(0, 1)
    f = _objective(sc, gt, torch.eye(4))

    # igs_track_lbfgs (gslam/frontend.py:604-662): 10 Adam steps, then L-BFGS, both at lr 0.002
    p = torch.nn.Parameter(torch.zeros(11))
    adam = torch.optim.Adam([p], 0.002)
    loss0 = None
    for _ in range(10):
        adam.zero_grad()
        loss = f(p)
        loss.backward()
        loss0 = float(loss) if loss0 is None else loss0
        adam.step()
    start = p.detach().clone()

    seen = []

    def evaluate(x):
        t = torch.tensor(np.asarray(x, dtype=np.float32), requires_grad=True)
        loss = f(t)
        loss.backward()
        seen.append(float(loss))
        return float(loss), t.grad.numpy().astype(np.float64)

    res = lbfgs_minimize(evaluate, start.numpy().astype(np.float64), lr=0.002, max_iter=20, history_size=5,
                         tolerance_change=1e-9)
    assert res.n_evals == len(seen) <= 26

    # torch's own optimizer from the same start on the same (deterministic, CPU) objective
    q = torch.nn.Parameter(start.clone())
    opt = torch.optim.LBFGS([q], history_size=5, line_search_fn="strong_wolfe", tolerance_change=1e-9, lr=0.002)
    last = []

    def closure():
        opt.zero_grad()
        loss = f(q)
        loss.backward()
        last.append(float(loss))
        return loss
    opt.step(closure)
    assert len(last) == res.n_evals
    assert abs(last[-1] - res.last_loss) <= 0.02 * last[-1], (last[-1], res.last_loss)
    assert np.allclose(res.x, q.detach().numpy(), atol=2e-3), np.abs(res.x - q.detach().numpy()).max()

    # and the schedule lowers the objective (pose recovery itself needs a realistic scene: the GPU
    # tests check it at 320x240 / 20k Gaussians)
    assert res.last_loss < 0.7 * loss0, (loss0, res.last_loss)
