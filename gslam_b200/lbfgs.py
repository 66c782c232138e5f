"""L-BFGS with a strong-Wolfe line search for a handful of unknowns, on the host.

The reference tracks every frame with ``torch.optim.LBFGS(params, history_size=5,
line_search_fn='strong_wolfe', tolerance_change=1e-9, lr=...)`` over 11 unknowns that live on
the GPU (gslam/frontend.py:604-662; the backend does the same for the window poses, gslam/
backend.py:436-506).  Every vector operation of that optimizer is then a CUDA launch on an
11-element tensor and every comparison a device->host sync; even with the unknowns moved to host
tensors the optimizer costs ~0.4 ms of tensor-op dispatch per objective evaluation, more than the
evaluation itself (0.28 ms for a 200k-Gaussian map).

This module restates the SAME algorithm -- PyTorch's ``LBFGS.step`` (one call: two-loop recursion
with `history_size` pairs, first step ``min(1, 1/|g|_1) * lr``, later steps ``lr``, stop rules on
max_iter / max_eval / tolerance_grad / tolerance_change) and its ``_strong_wolfe`` search
(bracketing with bounded cubic interpolation, zoom with the insufficient-progress rule, c1=1e-4,
c2=0.9, search iterations capped by the remaining evaluation budget) -- in NumPy float64 over a flat vector.  `evaluate(x)`
returns ``(loss, gradient)``; for tracking it is one CUDA-graph replay (gslam_b200.tracking).
The parity test (tests/test_lbfgs.py) runs both on the same float64 objectives and compares the
iterates and the number of evaluations.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import numpy as np

Evaluate = Callable[[np.ndarray], Tuple[float, np.ndarray]]


def _cubic_min(x1: float, f1: float, g1: float, x2: float, f2: float, g2: float,
               bounds: Optional[Tuple[float, float]] = None) -> float:
    """Minimiser of the cubic through two points with slopes, clipped to `bounds` (default: the
    interval between the points); the midpoint when the cubic has no real minimiser."""
    lo, hi = bounds if bounds is not None else ((x1, x2) if x1 <= x2 else (x2, x1))
    d1 = g1 + g2 - 3.0 * (f1 - f2) / (x1 - x2)
    disc = d1 * d1 - g1 * g2
    if disc < 0:
        return 0.5 * (lo + hi)
    d2 = math.sqrt(disc)
    if x1 <= x2:
        pos = x2 - (x2 - x1) * ((g2 + d2 - d1) / (g2 - g1 + 2.0 * d2))
    else:
        pos = x1 - (x1 - x2) * ((g1 + d2 - d1) / (g1 - g2 + 2.0 * d2))
    return min(max(pos, lo), hi)


@dataclass
class _Point:
    t: float
    f: float
    g: np.ndarray
    gtd: float


def strong_wolfe(phi: Callable[[float], Tuple[float, np.ndarray]], t: float, d: np.ndarray, f0: float,
                 g0: np.ndarray, gtd0: float, c1: float = 1e-4, c2: float = 0.9,
                 tolerance_change: float = 1e-9, max_ls: int = 25) -> Tuple[float, np.ndarray, float, int]:
    """Step length along `d` satisfying the strong Wolfe conditions.  `phi(t)` evaluates the objective
    at x + t d.  Returns (f, g, t, evaluations) of the accepted (or best bracketed) point."""
    d_norm = float(np.abs(d).max())
    f_new, g_new = phi(t)
    evals = 1
    new = _Point(t, f_new, g_new, float(g_new @ d))
    prev = _Point(0.0, f0, g0.copy(), gtd0)
    bracket: List[_Point] = []
    done = False
    it = 0
    while it < max_ls:
        if new.f > f0 + c1 * new.t * gtd0 or (it > 1 and new.f >= prev.f):
            bracket = [prev, new]
            break
        if abs(new.gtd) <= -c2 * gtd0:
            bracket = [new]
            done = True
            break
        if new.gtd >= 0:
            bracket = [prev, new]
            break
        # extrapolate: the next trial lies in [t + 0.01 (t - t_prev), 10 t]
        t_next = _cubic_min(prev.t, prev.f, prev.gtd, new.t, new.f, new.gtd,
                            bounds=(new.t + 0.01 * (new.t - prev.t), new.t * 10.0))
        prev = new
        f_new, g_new = phi(t_next)
        evals += 1
        new = _Point(t_next, f_new, g_new, float(g_new @ d))
        it += 1
    if it == max_ls:
        bracket = [_Point(0.0, f0, g0, gtd0), new]

    # zoom
    stalled = False
    lo, hi = (0, 1) if bracket[0].f <= bracket[-1].f else (1, 0)
    while not done and it < max_ls:
        a, b = bracket[0], bracket[1]
        if abs(b.t - a.t) * d_norm < tolerance_change:
            break
        t_try = _cubic_min(a.t, a.f, a.gtd, b.t, b.f, b.gtd)
        t_max, t_min = max(a.t, b.t), min(a.t, b.t)
        eps = 0.1 * (t_max - t_min)
        if min(t_max - t_try, t_try - t_min) < eps:
            if stalled or t_try >= t_max or t_try <= t_min:
                t_try = t_max - eps if abs(t_try - t_max) < abs(t_try - t_min) else t_min + eps
                stalled = False
            else:
                stalled = True
        else:
            stalled = False
        f_new, g_new = phi(t_try)
        evals += 1
        new = _Point(t_try, f_new, g_new, float(g_new @ d))
        it += 1
        if new.f > f0 + c1 * new.t * gtd0 or new.f >= bracket[lo].f:
            bracket[hi] = new
            lo, hi = (0, 1) if bracket[0].f <= bracket[1].f else (1, 0)
        else:
            if abs(new.gtd) <= -c2 * gtd0:
                done = True
            elif new.gtd * (bracket[hi].t - bracket[lo].t) >= 0:
                bracket[hi] = bracket[lo]
            bracket[lo] = new
    best = bracket[lo] if len(bracket) > 1 else bracket[0]
    return best.f, best.g, best.t, evals


@dataclass
class LbfgsResult:
    x: np.ndarray            # final unknowns (x_k + t d of the last accepted step)
    first_loss: float        # what torch's step() returns: the loss at the starting point
    last_loss: float         # loss of the LAST evaluation (the reference's `last_loss`, frontend.py:645)
    n_evals: int
    n_iters: int


def lbfgs_minimize(evaluate: Evaluate, x0: np.ndarray, lr: float = 1.0, max_iter: int = 20,
                   max_eval: Optional[int] = None, tolerance_grad: float = 1e-7,
                   tolerance_change: float = 1e-9, history_size: int = 100) -> LbfgsResult:
    """One ``torch.optim.LBFGS(..., line_search_fn='strong_wolfe').step(closure)`` on a flat vector."""
    if max_eval is None:
        max_eval = max_iter * 5 // 4
    x = np.asarray(x0, dtype=np.float64).copy()
    last = [float("nan")]

    def ev(z: np.ndarray) -> Tuple[float, np.ndarray]:
        f, g = evaluate(z)
        last[0] = float(f)
        return float(f), np.asarray(g, dtype=np.float64).copy()

    loss, grad = ev(x)
    first = loss
    evals = 1
    if float(np.abs(grad).max()) <= tolerance_grad:
        return LbfgsResult(x, first, last[0], evals, 0)
    ys: List[np.ndarray] = []      # gradient differences
    ss: List[np.ndarray] = []      # steps
    rho: List[float] = []
    h_diag = 1.0
    d = -grad
    t = 0.0
    prev_grad = grad
    n_iter = 0
    while n_iter < max_iter:
        n_iter += 1
        if n_iter > 1:
            y = grad - prev_grad
            s = d * t
            ys_dot = float(y @ s)
            if ys_dot > 1e-10:
                if len(ys) == history_size:
                    ys.pop(0), ss.pop(0), rho.pop(0)
                ys.append(y), ss.append(s), rho.append(1.0 / ys_dot)
                h_diag = ys_dot / float(y @ y)
            # two-loop recursion: d = -H grad
            q = -grad
            alpha = [0.0] * len(ys)
            for i in range(len(ys) - 1, -1, -1):
                alpha[i] = float(ss[i] @ q) * rho[i]
                q = q - alpha[i] * ys[i]
            d = q * h_diag
            for i in range(len(ys)):
                beta = float(ys[i] @ d) * rho[i]
                d = d + (alpha[i] - beta) * ss[i]
        prev_grad = grad
        prev_loss = loss
        t = min(1.0, 1.0 / float(np.abs(grad).sum())) * lr if n_iter == 1 else lr
        gtd = float(grad @ d)
        if gtd > -tolerance_change:
            break
        x_init = x
        # the search may spend at most what is left of the evaluation budget (torch passes max_ls likewise)
        loss, grad, t, ls_evals = strong_wolfe(lambda step: ev(x_init + step * d), t, d, loss, grad, gtd,
                                               max_ls=max_eval - evals)
        x = x_init + t * d
        evals += ls_evals
        if n_iter == max_iter or evals >= max_eval:
            break
        if float(np.abs(grad).max()) <= tolerance_grad:
            break
        if float(np.abs(d * t).max()) <= tolerance_change:
            break
        if abs(loss - prev_loss) < tolerance_change:
            break
    return LbfgsResult(x, first, last[0], evals, n_iter)


class DeviceLbfgs:
    """The same optimizer as `lbfgs_minimize`, resident on the GPU (csrc/lbfgs.cu): `reset(theta)` starts a
    minimisation from the device vector `theta`; after every evaluation of the objective at `theta` (loss and
    gradient in `result[1 + n]`, float32) `advance(theta, result)` writes the next point to evaluate into
    `theta`.  Nothing here synchronises: the caller replays evaluate + advance `max_eval` times (one CUDA
    graph) and reads `summary` (finished, evaluations used, iterations, first loss, last loss) once."""

    def __init__(self, n: int, device, lr: float = 1.0, max_iter: int = 20, max_eval: Optional[int] = None,
                 tolerance_grad: float = 1e-7, tolerance_change: float = 1e-9, history_size: int = 100):
        import torch
        from ._lib import lib
        self.n, self.lr, self.max_iter = int(n), float(lr), int(max_iter)
        self.max_eval = int(max_eval) if max_eval is not None else max_iter * 5 // 4
        self.tol_grad, self.tol_change = float(tolerance_grad), float(tolerance_change)
        self.hist = min(int(history_size), 16)
        if n > 96:
            raise ValueError("DeviceLbfgs holds at most 96 unknowns")
        self.state = torch.zeros(lib().gsb_lbfgs_state_bytes(), dtype=torch.uint8, device=device)
        self.summary = torch.zeros(5, dtype=torch.float32, device=device)

    def reset(self, theta) -> None:
        from ._lib import call, ptr, stream
        call("gsb_lbfgs_init", ptr(self.state), ptr(theta), self.n, self.hist, self.max_iter, self.max_eval, self.lr,
             self.tol_grad, self.tol_change, stream())

    def advance(self, theta, result) -> None:
        from ._lib import call, ptr, stream
        call("gsb_lbfgs_advance", ptr(self.state), ptr(theta), ptr(result), ptr(self.summary), stream())
