"""gslam_b200.lbfgs against torch.optim.LBFGS (the optimizer the reference calls,
gslam/frontend.py:614-620, gslam/backend.py:455-460) on the same float64 objectives: same iterates,
same number of objective evaluations, same reported losses."""
import numpy as np
import pytest
import torch

from gslam_b200.lbfgs import lbfgs_minimize


def _rosenbrock(x):
    return (100.0 * (x[1:] - x[:-1] ** 2) ** 2 + (1 - x[:-1]) ** 2).sum()


def _quadratic(x):
    n = x.shape[0]
    A = torch.diag(torch.linspace(1.0, 50.0, n, dtype=x.dtype)) + 0.3
    b = torch.linspace(-1.0, 2.0, n, dtype=x.dtype)
    return 0.5 * x @ A @ x - b @ x


def _bumpy(x):
    return (x ** 2).sum() + 0.5 * torch.sin(5 * x).sum() + 0.1 * torch.cos(3 * x[0] * x[1])


def _photometric_like(x):
    # small gradients and a loss near 1e-3, like the tracking objective
    return 1e-3 * (torch.tanh(x * 3 - 0.2) ** 2).sum() + 1e-4 * (x ** 4).sum()


CASES = [(_rosenbrock, 11, 1.0, 5), (_rosenbrock, 11, 0.002, 5), (_quadratic, 11, 1.0, 5),
         (_quadratic, 9, 0.002, 10), (_bumpy, 11, 1.0, 5), (_bumpy, 11, 0.002, 5),
         (_photometric_like, 11, 0.002, 5), (_photometric_like, 11, 1.0, 3)]


@pytest.mark.parametrize("fn,n,lr,hist", CASES)
def test_same_iterates_and_evaluation_count_as_torch(fn, n, lr, hist):
    g = torch.Generator().manual_seed(n + hist)
    x0 = (torch.rand(n, generator=g, dtype=torch.float64) - 0.5) * 0.8

    p = torch.nn.Parameter(x0.clone())
    opt = torch.optim.LBFGS([p], history_size=hist, line_search_fn="strong_wolfe", tolerance_change=1e-9, lr=lr)
    seen = []

    def closure():
        opt.zero_grad()
        loss = fn(p)
        loss.backward()
        seen.append(float(loss))
        return loss
    first = float(opt.step(closure))

    ours_seen = []

    def evaluate(z):
        t = torch.tensor(z, dtype=torch.float64, requires_grad=True)
        loss = fn(t)
        loss.backward()
        ours_seen.append(float(loss))
        return float(loss), t.grad.numpy()
    r = lbfgs_minimize(evaluate, x0.numpy(), lr=lr, max_iter=20, history_size=hist, tolerance_change=1e-9)

    assert r.n_evals == len(seen) == len(ours_seen)
    assert np.allclose(ours_seen, seen, rtol=1e-9, atol=1e-12)
    assert np.allclose(r.x, p.detach().numpy(), rtol=1e-8, atol=1e-10)
    assert r.first_loss == pytest.approx(first, rel=1e-12)
    assert r.last_loss == pytest.approx(seen[-1], rel=1e-9)


def test_stops_at_once_on_a_flat_gradient():
    r = lbfgs_minimize(lambda z: (1.0, np.zeros_like(z)), np.ones(4))
    assert r.n_evals == 1 and r.n_iters == 0 and np.array_equal(r.x, np.ones(4))


def test_respects_max_iter_and_max_eval():
    def evaluate(z):
        t = torch.tensor(z, dtype=torch.float64, requires_grad=True)
        loss = _rosenbrock(t)
        loss.backward()
        return float(loss), t.grad.numpy()
    x0 = np.full(6, -0.7)
    r = lbfgs_minimize(evaluate, x0, lr=1.0, max_iter=3)
    assert r.n_iters <= 3
    p = torch.nn.Parameter(torch.tensor(x0))
    opt = torch.optim.LBFGS([p], max_iter=3, line_search_fn="strong_wolfe")

    def closure():
        opt.zero_grad()
        loss = _rosenbrock(p)
        loss.backward()
        return loss
    opt.step(closure)
    assert np.allclose(r.x, p.detach().numpy(), rtol=1e-8, atol=1e-10)
