"""Frame-to-map tracking on a frozen map, the callers of the render path in
the reference frontend:

* ``Frontend.track`` (gslam/frontend.py:170-247): constant-motion initial pose, fresh exposure
  parameters, then the pose optimisation below;
* ``Frontend.igs_track_lbfgs`` (gslam/frontend.py:604-662): 10 Adam steps, then
  ``torch.optim.LBFGS(history_size=5, line_search_fn='strong_wolfe', tolerance_change=1e-9)``;
* ``Frontend.igs_track`` (gslam/frontend.py:573-602): SGD(momentum 0.8, nesterov) with an
  exponential learning-rate decay and ``StopOnPlateau(20, 0.1)`` (gslam/utils.py:164-186).

The unknowns of one frame are 11 floats -- ``PoseZhou.dR[6]``, ``PoseZhou.dt[3]`` and the exposure
``(a, b)`` -- and every evaluation of the objective is one render + loss + backward over the whole
map.  The reference spends ~25 autograd nodes on the pose chain, ~12 on the loss and a ``.item()``
per evaluation; its L-BFGS runs dozens of 11-element vector kernels per iteration on the GPU with
host syncs in the line search.  Here:

* the evaluation (pose chain -> ``rasterization()`` -> tracking loss -> backward to the 11
  unknowns) is captured ONCE per map as a CUDA graph and replayed; the frame's image, its initial
  pose and the unknowns live in static device buffers, so the same graph serves every frame until
  the map changes (``set_map``);
* the Adam / SGD phases replay a graph that also contains the optimizer update, with no host sync
  at all (the loss history is written to a device ring and read once);
* L-BFGS keeps the 11 unknowns in pinned host memory and runs `gslam_b200.lbfgs` on them, a NumPy
  restatement of ``torch.optim.LBFGS`` + strong Wolfe (the reference's optimizer, same options;
  checked iterate-for-iterate against torch in tests/test_lbfgs.py): each evaluation is one graph
  launch that starts with the 44-byte upload of the unknowns and ends with the 48-byte download of
  (loss, gradient), and ONE stream sync -- the two-loop recursion and the cubic interpolation are
  11-element host arithmetic, not GPU launches (torch's optimizer on host tensors costs ~0.4 ms of
  op dispatch per evaluation, more than the 0.28 ms evaluation itself).

CUDA only: there is no CPU path (the closure is the sm_100a render path).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional

import torch
from torch import Tensor

from . import ops
from .graph import GraphedStep
from .lbfgs import DeviceLbfgs, lbfgs_minimize
from .losses import tracking_loss
from .map import GaussianSplattingData
from .primitives import Camera, pose_chain, affine_inverse  # noqa: F401  (re-exported)

N_POSE = 9          # dR[6] | dt[3]
N_THETA = 11        # ... | exposure (a, b)


@dataclass
class TrackingConfig:
    """The fields of the reference TrackingConfig (gslam/frontend.py:45-61) that tracking reads."""
    num_tracking_iters: int = 200
    photometric_loss: str = "active-nerf"      # 'l1' | 'mse' | 'active-nerf'
    pose_optim_lr: float = 0.002
    pose_optim_lr_decay: float = 0.99
    learn_exposure_params: bool = True
    use_gt_depths: bool = False
    n_adam_iters: int = 10                     # the literal 10 of igs_track_lbfgs (:649)
    lbfgs_history: int = 5
    lbfgs_max_iter: int = 20                   # torch.optim.LBFGS default (max_eval = 25); 0 skips the phase
    lbfgs_tolerance_change: float = 1e-9
    exposure_lr: float = 0.01                  # SGD param-group lr of the exposure (:204-209)
    lbfgs_on_device: bool = True               # the L-BFGS state machine runs on the GPU (csrc/lbfgs.cu): one sync per
                                               # frame; False: gslam_b200.lbfgs on the host, one sync per evaluation
    deterministic: bool = False                # fixed-order gradient accumulation (ops.deterministic): the same
                                               # frame gives bit-identical poses from run to run


class StopOnPlateau:
    """gslam/utils.py:164-186, same rule."""

    def __init__(self, patience: int, min_loss: float):
        self.patience, self.min_loss = patience, min_loss
        self.counter, self.last_loss = 0, None

    def stop(self, loss: float) -> bool:
        if self.last_loss is None:
            self.last_loss = loss
            return False
        if loss > self.min_loss:
            return False
        if self.last_loss > loss:
            self.counter += 1
            if self.counter >= self.patience:
                return True
        else:
            self.counter = 0
        self.last_loss = loss
        return False


class _SplitTheta(torch.autograd.Function):
    """theta[11] -> (dR[1,6], dt[1,3], exposure[1,2]); the backward is ONE cat (autograd's three
    slice nodes would each zero-fill and accumulate an 11-vector)."""

    @staticmethod
    def forward(ctx, theta: Tensor):
        return theta[0:6].view(1, 6), theta[6:9].view(1, 3), theta[9:11].view(1, 2)

    @staticmethod
    def backward(ctx, g_dR, g_dt, g_ex):
        dev = next(g for g in (g_dR, g_dt, g_ex) if g is not None).device
        z = lambda n: torch.zeros(n, device=dev)  # noqa: E731
        return torch.cat([g_dR.reshape(6) if g_dR is not None else z(6),
                          g_dt.reshape(3) if g_dt is not None else z(3),
                          g_ex.reshape(2) if g_ex is not None else z(2)])


@dataclass
class TrackResult:
    viewmat: Tensor            # [4,4] optimised world -> camera matrix (device)
    exposure: Tensor           # [2] (device)
    loss: float
    n_evals: int               # objective evaluations (graph replays)
    n_syncs: int               # host <-> device synchronisations spent on them


class Tracker:
    """Pose + exposure optimisation of single frames against a frozen map."""

    def __init__(self, camera: Camera, conf: Optional[TrackingConfig] = None, device="cuda"):
        self.conf = conf or TrackingConfig()
        self.camera = camera
        self.dev = torch.device(device)
        if self.dev.type != "cuda":
            raise RuntimeError("gslam_b200.tracking.Tracker needs a CUDA device (no CPU fallback)")
        H, W = camera.height, camera.width
        d = self.dev
        # static buffers every captured graph reads / writes
        self.Rt = torch.eye(4, device=d).view(1, 4, 4).contiguous()
        self.gt = torch.zeros(1, H, W, 3, device=d)
        self.gt_depth = torch.zeros(1, H, W, device=d) if self.conf.use_gt_depths else None
        self.theta = torch.zeros(N_THETA, device=d, requires_grad=True)
        self.result = torch.zeros(1 + N_THETA, device=d)              # loss | gradient
        self.theta_host = torch.zeros(N_THETA).pin_memory()
        self.result_host = torch.zeros(1 + N_THETA).pin_memory()
        # optimizer state of the sync-free phases (Adam moments / SGD momentum, step, lr scale)
        self.m = torch.zeros(N_THETA, device=d)
        self.v = torch.zeros(N_THETA, device=d)
        self.step_no = torch.zeros((), device=d)
        self.lr_scale = torch.ones((), device=d)
        self.loss_ring = torch.zeros(1024, device=d)
        self.ring_pos = torch.zeros((), dtype=torch.long, device=d)
        mask = torch.ones(N_THETA, device=d)
        if not self.conf.learn_exposure_params:
            mask[N_POSE:] = 0
        self.mask = mask
        lr_sgd = torch.full((N_THETA,), self.conf.pose_optim_lr, device=d)
        lr_sgd[N_POSE:] = self.conf.exposure_lr
        self.lr_sgd = lr_sgd * mask
        self.splats: Optional[GaussianSplattingData] = None
        self.capacity = 0
        self.n_live = 0
        self.n_captures = 0          # graph captures so far (2 per map capacity, more only after an overflow)
        self._graphs = {}
        self.headroom = 1.5          # capacity of the captured key buffers over the measured count
        self.n_evals = 0
        self.n_syncs = 0
        c = self.conf
        self.dev_lbfgs = DeviceLbfgs(N_THETA, d, lr=c.pose_optim_lr, max_iter=max(c.lbfgs_max_iter, 1),
                                     tolerance_change=c.lbfgs_tolerance_change, history_size=c.lbfgs_history)
        self.summary_host = torch.zeros(5 + N_THETA).pin_memory()
        self.lbfgs_evals_used = 0

    # ---- the map -------------------------------------------------------------------------------
    def set_map(self, splats: GaussianSplattingData) -> None:
        """Frontend.sync (gslam/frontend.py:254-269): a frozen copy of the backend's map.

        The captured graphs bake in the parameter pointers and the Gaussian count, and the map grows
        and shrinks between syncs.  So the copy goes into buffers of fixed CAPACITY (1.25 x the count,
        rounded up to 1024 rows): rows [0, N) hold the map, the rest are "dead" rows parked 1e15 m
        away, which the projection culls against the far plane (radius 0, no tile, no pixel), so the
        render is that of the N-row map.  A sync is then one device-to-device copy and the graphs
        survive it; only a map that outgrows the capacity allocates anew and drops them."""
        src = [getattr(splats, name).detach() for name in GaussianSplattingData._per_splat_params[:6]]
        n = src[0].shape[0]
        if self.splats is None or n > self.capacity:
            self.capacity = max(1024, -(-int(n * 1.25) // 1024) * 1024)
            d, cap = self.dev, self.capacity
            self.splats = GaussianSplattingData(
                torch.empty(cap, 3, device=d), torch.empty(cap, 4, device=d), torch.empty(cap, 3, device=d),
                torch.empty(cap, device=d), torch.empty(cap, 3, device=d), torch.empty(cap, device=d),
                torch.zeros(cap, dtype=torch.long, device=d)).to(d)
            for name in GaussianSplattingData._per_splat_params:
                getattr(self.splats, name).requires_grad_(False)
            self._graphs = {}
        dst = [getattr(self.splats, name) for name in GaussianSplattingData._per_splat_params[:6]]
        dead = (1e15, None, math.log(1e-3), -20.0, 0.0, 0.0)
        with torch.no_grad():
            for k, (t, v) in enumerate(zip(dst, dead)):
                t[:n].copy_(src[k])
                if k == 1:
                    t[n:] = t.new_tensor([1.0, 0.0, 0.0, 0.0])
                else:
                    t[n:] = v
        self.n_live = n

    # ---- the objective -------------------------------------------------------------------------
    def _evaluate(self) -> Tensor:
        """loss and d loss / d theta of the current device-side unknowns -> self.result."""
        self.theta.grad = None
        dR, dt, ex = _SplitTheta.apply(self.theta)
        viewmats = pose_chain(self.Rt, dR, dt)
        out = self.splats._rasterize([self.camera], viewmats, "RGB+D", 0.5)     # = splats.render(...), gslam/map.py:50-76
        exposure = ex if self.conf.learn_exposure_params else None
        # backward() below starts from this loss (plus, at most, an added depth term): coefficient 1
        loss = tracking_loss(out.render, self.gt, exposure, self.conf.photometric_loss, unit_upstream=True)
        if self.conf.use_gt_depths:
            # (rendered - gt)[gt > 0].abs().mean() * 0.01 without the data-dependent shape
            valid = (self.gt_depth > 0).float()
            loss = loss + 0.01 * ((out.depthmaps - self.gt_depth).abs() * valid).sum() / valid.sum().clamp_min(1.0)
        loss.backward()
        torch.cat([loss.detach().view(1), self.theta.grad * self.mask], out=self.result)
        return loss

    def _log_loss(self) -> None:
        self.loss_ring.index_copy_(0, self.ring_pos.view(1) % self.loss_ring.shape[0], self.result[0:1])
        self.ring_pos += 1

    def _step_closure(self) -> Tensor:
        """L-BFGS closure body: unknowns up, evaluate, (loss, gradient) down."""
        with torch.no_grad():
            self.theta.copy_(self.theta_host, non_blocking=True)
        loss = self._evaluate()
        self.result_host.copy_(self.result, non_blocking=True)
        return loss

    def _step_closure_dev(self) -> Tensor:
        """L-BFGS evaluation with the optimizer on the device: evaluate at theta, let the state machine
        consume (loss, gradient) and write the next trial point into theta.  No transfer, no sync."""
        loss = self._evaluate()
        self.dev_lbfgs.advance(self.theta, self.result)
        return loss

    def _step_adam(self) -> Tensor:
        """torch.optim.Adam(params, lr) defaults (betas .9/.999, eps 1e-8) on the 11 unknowns."""
        loss = self._evaluate()
        with torch.no_grad():
            g = self.result[1:]
            self.step_no += 1
            self.m.lerp_(g, 0.1)
            self.v.mul_(0.999).addcmul_(g, g, value=0.001)
            bc1 = 1 - torch.pow(0.9, self.step_no)
            bc2 = 1 - torch.pow(0.999, self.step_no)
            denom = (self.v / bc2).sqrt_().add_(1e-8)
            self.theta.sub_(self.conf.pose_optim_lr * (self.m / bc1) / denom * self.mask)
            self._log_loss()
        return loss

    def _step_sgd(self) -> Tensor:
        """torch.optim.SGD(momentum=0.8, nesterov=True) + ExponentialLR (gslam/frontend.py:194-209)."""
        loss = self._evaluate()
        with torch.no_grad():
            g = self.result[1:]
            first = (self.step_no == 0).float()
            self.m.copy_(first * g + (1 - first) * (0.8 * self.m + g))
            self.theta.sub_(self.lr_sgd * self.lr_scale * (g + 0.8 * self.m))
            self.step_no += 1
            self.lr_scale *= self.conf.pose_optim_lr_decay
            self._log_loss()
        return loss

    def _graph(self, kind: str) -> GraphedStep:
        g = self._graphs.get(kind)
        if g is None:
            fn = {"closure": self._step_closure, "closure_dev": self._step_closure_dev, "adam": self._step_adam,
                  "sgd": self._step_sgd}[kind]
            saved = [t.clone() for t in (self.theta.detach(), self.m, self.v, self.step_no, self.lr_scale,
                                         self.ring_pos)]
            with ops.deterministic_mode(self.conf.deterministic):     # baked into the captured kernels
                g = GraphedStep(fn, headroom=self.headroom)
            self.n_captures += 1
            # the eager warm-up runs of the capture moved the unknowns; put them back
            with torch.no_grad():
                for t, s in zip((self.theta, self.m, self.v, self.step_no, self.lr_scale, self.ring_pos), saved):
                    t.copy_(s)
            self._graphs[kind] = g
        return g

    # ---- one frame ------------------------------------------------------------------------------
    def _begin(self, img: Tensor, init_viewmat: Tensor, exposure_init: Optional[Tensor],
               gt_depth: Optional[Tensor]) -> None:
        if self.splats is None:
            raise RuntimeError("Tracker.set_map() has not been called")
        with torch.no_grad():
            self.gt[0].copy_(img)
            if self.gt_depth is not None and gt_depth is not None:
                self.gt_depth[0].copy_(gt_depth)
            self.Rt[0].copy_(init_viewmat)
            self.theta.zero_()
            if exposure_init is not None and self.conf.learn_exposure_params:
                self.theta[N_POSE:].copy_(exposure_init)
            for t in (self.m, self.v, self.step_no, self.ring_pos):
                t.zero_()
            self.lr_scale.fill_(1.0)
        for g in self._graphs.values():
            g.reset_overflow()
        self.n_evals = 0
        self.n_syncs = 0

    def _finish(self, loss: float) -> TrackResult:
        with torch.no_grad():
            th = self.theta.detach()
            viewmat = pose_chain(self.Rt, th[0:6].view(1, 6), th[6:9].view(1, 3))[0].clone()
            exposure = th[N_POSE:].clone()
        return TrackResult(viewmat, exposure, loss, self.n_evals, self.n_syncs)

    def _losses(self, n: int) -> List[float]:
        self.n_syncs += 1
        return self.loss_ring[:n].tolist()

    def _track_lbfgs(self, img: Tensor, init_viewmat: Tensor, exposure_init: Optional[Tensor] = None,
                    gt_depth: Optional[Tensor] = None) -> TrackResult:
        """igs_track_lbfgs (gslam/frontend.py:604-662)."""
        self._begin(img, init_viewmat, exposure_init, gt_depth)
        c = self.conf
        if c.n_adam_iters:
            adam = self._graph("adam")
            for _ in range(c.n_adam_iters):
                adam.replay()
            self.n_evals += c.n_adam_iters
        if c.lbfgs_max_iter <= 0:
            loss = self._losses(c.n_adam_iters)[-1] if c.n_adam_iters else float("nan")
            adam.check() if c.n_adam_iters else None
            return self._finish(loss)
        if c.lbfgs_on_device:
            return self._lbfgs_on_device(adam if c.n_adam_iters else None)
        closure_graph = self._graph("closure")
        self.theta_host.copy_(self.theta.detach())            # syncs: the Adam phase is done
        self.n_syncs += 1
        if c.n_adam_iters:
            adam.check()
        theta_np, result_np = self.theta_host.numpy(), self.result_host.numpy()    # views of the pinned buffers
        stream = torch.cuda.current_stream(self.dev)

        def evaluate(x):
            theta_np[:] = x                                   # float64 -> the float32 the kernels see
            closure_graph.replay()                            # 44 B up, evaluation, 48 B down
            stream.synchronize()
            self.n_evals += 1
            self.n_syncs += 1
            return float(result_np[0]), result_np[1:]

        res = lbfgs_minimize(evaluate, theta_np.copy(), lr=c.pose_optim_lr, max_iter=c.lbfgs_max_iter,
                             tolerance_change=c.lbfgs_tolerance_change, history_size=c.lbfgs_history)
        theta_np[:] = res.x
        last = [res.last_loss]
        closure_graph.check()
        with torch.no_grad():
            self.theta.copy_(self.theta_host)
        return self._finish(last[0])

    def _lbfgs_on_device(self, adam: Optional[GraphedStep]) -> TrackResult:
        """The L-BFGS phase with the optimizer on the GPU: max_eval replays of (evaluate, advance) queued back
        to back, ONE read-back at the end (the reference syncs inside every line-search comparison)."""
        g = self._graph("closure_dev")        # (its capture warm-up moved theta and the optimizer state: reset after)
        self.dev_lbfgs.reset(self.theta)
        # a search may spend one evaluation more than the budget left (its first trial is not counted against
        # max_ls, as in torch): max_eval + 1 bounds the whole minimisation
        n = self.dev_lbfgs.max_eval + 1
        for _ in range(n):
            g.replay()
        self.n_evals += n
        with torch.no_grad():
            self.summary_host.copy_(torch.cat([self.dev_lbfgs.summary, self.theta.detach()]), non_blocking=True)
        torch.cuda.current_stream(self.dev).synchronize()
        self.n_syncs += 1
        if adam is not None:
            adam.check()
        g.check()
        self.lbfgs_evals_used = int(self.summary_host[1])
        return self._finish(float(self.summary_host[4]))

    def _track_sgd(self, img: Tensor, init_viewmat: Tensor, exposure_init: Optional[Tensor] = None,
                  gt_depth: Optional[Tensor] = None, n_iters: Optional[int] = None,
                  check_every: int = 20) -> TrackResult:
        """igs_track (gslam/frontend.py:573-602).  The reference reads the loss back every iteration
        for StopOnPlateau(20, 0.1); here the iterations run in sync-free chunks of `check_every`
        and the same rule is applied to the chunk's loss history, so a stop can come up to
        check_every - 1 iterations later than the reference's."""
        self._begin(img, init_viewmat, exposure_init, gt_depth)
        n_iters = self.conf.num_tracking_iters if n_iters is None else n_iters
        assert n_iters <= self.loss_ring.shape[0]
        sgd = self._graph("sgd")
        stop = StopOnPlateau(20, 0.1)
        done, loss, stopped = 0, float("nan"), False
        while done < n_iters and not stopped:
            chunk = min(check_every, n_iters - done)
            for _ in range(chunk):
                sgd.replay()
            self.n_evals += chunk
            losses = self._losses(done + chunk)
            for k in range(done, done + chunk):
                loss = losses[k]
                if stop.stop(loss):
                    stopped = True
                    break
            done += chunk
        sgd.check()
        return self._finish(loss)


    def _retrying(self, fn, *args, **kwargs) -> TrackResult:
        """A frame whose pose sees more tile intersections than the captured key buffers hold
        reports an overflow (graph.GraphedStep.check); re-capture with twice the headroom, redo."""
        for attempt in range(3):
            try:
                return fn(*args, **kwargs)
            except RuntimeError as e:
                if "exceed the captured capacity" not in str(e) or attempt == 2:
                    raise
                self.headroom *= 2.0
                self._graphs = {}

    def track_lbfgs(self, img: Tensor, init_viewmat: Tensor, exposure_init: Optional[Tensor] = None,
                    gt_depth: Optional[Tensor] = None) -> TrackResult:
        """igs_track_lbfgs (gslam/frontend.py:604-662)."""
        return self._retrying(self._track_lbfgs, img, init_viewmat, exposure_init, gt_depth)

    def track_sgd(self, img: Tensor, init_viewmat: Tensor, exposure_init: Optional[Tensor] = None,
                  gt_depth: Optional[Tensor] = None, n_iters: Optional[int] = None,
                  check_every: int = 20) -> TrackResult:
        """igs_track (gslam/frontend.py:573-602); see _track_sgd for the stop rule."""
        return self._retrying(self._track_sgd, img, init_viewmat, exposure_init, gt_depth, n_iters, check_every)


def constant_motion(prev2: Optional[Tensor], prev1: Tensor) -> Tensor:
    """Initial pose of a new frame (gslam/frontend.py:176-183): pose_b @ inv(pose_a) @ pose_b."""
    if prev2 is None:
        return prev1.clone()
    return prev1 @ affine_inverse(prev2) @ prev1
