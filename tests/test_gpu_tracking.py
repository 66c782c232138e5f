"""Tracking on a frozen map: the graph-replayed optimisation phases of
gslam_b200.tracking against the reference's own loop structure (gslam/frontend.py:573-662) --
PoseZhou module + exposure parameter + stock torch optimizers, stepped eagerly through the
same render path."""
import pytest
import torch

from gslam_b200.losses import tracking_loss
from gslam_b200.map import GaussianSplattingData
from gslam_b200.primitives import Camera, PoseZhou
from gslam_b200.synthetic import make_scene, smooth_trajectory
from gslam_b200.tracking import Tracker, TrackingConfig

pytestmark = pytest.mark.gpu


def _problem(cuda, n=20_000, frame=10):
    sc = make_scene(n, 1, "tum", 0.5).to(cuda)
    splats = GaussianSplattingData(*sc.params(), torch.zeros(n, dtype=torch.long, device=cuda)).no_grad_clone()
    cam = Camera(sc.Ks[0].clone(), sc.height, sc.width)
    true = smooth_trajectory(frame + 1)[frame].to(cuda)
    with torch.no_grad():
        gt = splats.render([cam], [true]).rgbs[0].clamp(0, 1).clone()
    return splats, cam, gt, true


def _reference_closure(splats, cam, pose, exposure, gt):
    def closure():
        out = splats([cam], [pose], render_depth=True)
        loss = tracking_loss(out.render, gt[None], exposure, "active-nerf")
        loss.backward()
        return loss
    return closure


def _pose_error(a, b):
    return float((torch.linalg.inv(a) @ b - torch.eye(4, device=a.device)).norm())


def test_adam_phase_matches_torch_adam_on_the_pose_module(cuda):
    splats, cam, gt, _ = _problem(cuda)
    pose = PoseZhou(torch.eye(4, device=cuda)).to(cuda)
    exposure = torch.zeros(2, device=cuda, requires_grad=True)
    opt = torch.optim.Adam([pose.dR, pose.dt, exposure], 0.002)
    closure = _reference_closure(splats, cam, pose, exposure, gt)
    for _ in range(10):
        loss_ref = closure()
        opt.step()
        opt.zero_grad()

    tr = Tracker(cam, TrackingConfig(lbfgs_max_iter=0), cuda)
    tr.set_map(splats)
    r = tr.track_lbfgs(gt, torch.eye(4, device=cuda))
    assert r.n_evals == 10 and r.n_syncs <= 2
    assert torch.allclose(r.viewmat, pose().detach(), atol=3e-5), (r.viewmat - pose().detach()).abs().max()
    assert torch.allclose(r.exposure, exposure.detach(), atol=3e-5)
    assert abs(r.loss - float(loss_ref)) <= 1e-3 * abs(float(loss_ref)) + 1e-7


def test_sgd_phase_matches_torch_sgd_nesterov_with_exponential_decay(cuda):
    splats, cam, gt, _ = _problem(cuda)
    pose = PoseZhou(torch.eye(4, device=cuda)).to(cuda)
    exposure = torch.zeros(2, device=cuda, requires_grad=True)
    opt = torch.optim.SGD(pose.parameters(), 0.002, momentum=0.8, nesterov=True)
    sched = torch.optim.lr_scheduler.ExponentialLR(opt, 0.99)
    opt.add_param_group({"params": exposure, "lr": 0.01})
    closure = _reference_closure(splats, cam, pose, exposure, gt)
    for _ in range(12):
        closure()
        opt.step()
        sched.step()
        opt.zero_grad()

    tr = Tracker(cam, TrackingConfig(), cuda)
    tr.set_map(splats)
    r = tr.track_sgd(gt, torch.eye(4, device=cuda), n_iters=12, check_every=6)
    assert r.n_evals == 12 and r.n_syncs == 2
    assert torch.allclose(r.viewmat, pose().detach(), atol=3e-5), (r.viewmat - pose().detach()).abs().max()
    assert torch.allclose(r.exposure, exposure.detach(), atol=3e-5)


def test_lbfgs_tracking_recovers_the_pose_like_the_reference_loop(cuda):
    splats, cam, gt, true = _problem(cuda)
    # the reference's igs_track_lbfgs, eagerly, everything on the device
    pose = PoseZhou(torch.eye(4, device=cuda)).to(cuda)
    exposure = torch.zeros(2, device=cuda, requires_grad=True)
    params = [pose.dR, pose.dt, exposure]
    closure = _reference_closure(splats, cam, pose, exposure, gt)
    adam = torch.optim.Adam(params, 0.002)
    loss0 = None
    for _ in range(10):
        l = closure()
        loss0 = float(l) if loss0 is None else loss0
        adam.step()
        adam.zero_grad()
    lbfgs = torch.optim.LBFGS(params, history_size=5, line_search_fn="strong_wolfe", tolerance_change=1e-9, lr=0.002)
    last = [None]

    def lbfgs_closure():
        lbfgs.zero_grad()
        l = closure()
        last[0] = float(l)
        return l
    lbfgs.step(lbfgs_closure)

    tr = Tracker(cam, TrackingConfig(), cuda)
    tr.set_map(splats)
    r = tr.track_lbfgs(gt, torch.eye(4, device=cuda))
    start = _pose_error(torch.eye(4, device=cuda), true)
    err, err_ref = _pose_error(r.viewmat, true), _pose_error(pose().detach(), true)
    assert r.loss < 0.5 * loss0, (r.loss, loss0)
    # Same algorithm and schedule as the torch loop above; the claim is relative to THAT loop run on the
    # same problem (what the schedule -- lr 0.002, 10 Adam + <= 25 L-BFGS evaluations, free exposure --
    # can recover is a property of the problem, not of the tracker): float64 host L-BFGS vs torch's
    # float32 one and reordered atomics perturb the line search, so close, not equal.
    assert r.loss <= 1.5 * last[0] + 1e-7, (r.loss, last[0])
    assert err < start, (err, err_ref, start)
    assert err <= 1.5 * err_ref + 1e-3, (err, err_ref, start)
    # one sync per L-BFGS evaluation, none in the Adam phase
    assert r.n_syncs <= r.n_evals - 10 + 1


def test_graphs_are_captured_once_per_map_and_reused_across_frames(cuda):
    splats, cam, gt, true = _problem(cuda, n=8_000)
    tr = Tracker(cam, TrackingConfig(deterministic=True), cuda)
    tr.set_map(splats)
    a = tr.track_lbfgs(gt, torch.eye(4, device=cuda))
    graphs = dict(tr._graphs)
    b = tr.track_lbfgs(gt.flip(1), a.viewmat)           # another image, another start
    assert all(tr._graphs[k] is g for k, g in graphs.items()) and len(tr._graphs) == 2
    c = tr.track_lbfgs(gt, torch.eye(4, device=cuda))   # the first problem again: same answer
    # bitwise, because the tracker runs in deterministic mode (with atomics the gradient's last bits
    # change between replays and the line search amplifies that into another path)
    assert torch.equal(a.viewmat, c.viewmat) and a.loss == c.loss
    assert torch.isfinite(b.viewmat).all()
    # a sync with a map that fits the capacity keeps the graphs; dead rows render nothing
    assert tr.capacity >= 8_000 and tr.n_live == 8_000
    tr.set_map(splats.mask(torch.arange(8_000, device=cuda) < 7_000))
    assert tr.n_live == 7_000 and all(tr._graphs[k] is g for k, g in graphs.items())
    d = tr.track_lbfgs(gt, torch.eye(4, device=cuda))
    assert torch.isfinite(d.viewmat).all() and d.loss < 10 * a.loss
    out = tr.splats.render([cam], [true])
    assert int((out.radii[0, 7_000:] != 0).sum()) == 0
    # outgrowing the capacity allocates anew and drops them
    big, _, _, _ = _problem(cuda, n=tr.capacity + 1)
    tr.set_map(big)
    assert not tr._graphs and tr.capacity > 8_000 * 1.25


def test_frozen_exposure_stays_zero(cuda):
    splats, cam, gt, _ = _problem(cuda, n=8_000)
    tr = Tracker(cam, TrackingConfig(learn_exposure_params=False, lbfgs_max_iter=4), cuda)
    tr.set_map(splats)
    r = tr.track_lbfgs(gt * 0.8 + 0.05, torch.eye(4, device=cuda))
    assert torch.equal(r.exposure, torch.zeros(2, device=cuda))
