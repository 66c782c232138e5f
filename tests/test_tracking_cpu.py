"""Host-side pieces of the tracking / sequence loop (no GPU): the stop rule, the constant-motion
start, the packed-unknowns split, and the loud failure without CUDA."""
import pytest
import torch

from gslam_b200.synthetic import smooth_trajectory
from gslam_b200.tracking import StopOnPlateau, Tracker, _SplitTheta, constant_motion, affine_inverse
from gslam_b200.primitives import Camera


def test_stop_on_plateau_follows_the_reference_rule():
    # gslam/utils.py:164-186: counts DEcreases below min_loss, resets on a non-decrease
    s = StopOnPlateau(3, 0.1)
    assert not s.stop(1.0)                       # first call only records
    assert not s.stop(0.5)                       # above min_loss: ignored, last_loss stays 1.0
    assert not s.stop(0.09) and s.counter == 1
    assert not s.stop(0.08) and s.counter == 2
    assert not s.stop(0.085) and s.counter == 0  # went up: reset
    assert not s.stop(0.07)
    assert not s.stop(0.06)
    assert s.stop(0.05)


def test_constant_motion_extrapolates_a_rigid_step():
    step = smooth_trajectory(3)[1]
    a = torch.eye(4)
    b = step @ a
    c = constant_motion(a, b)
    assert torch.allclose(c, step @ step, atol=1e-6)
    assert torch.equal(constant_motion(None, b), b)


def test_split_theta_gradient_is_the_concatenation():
    theta = torch.arange(11.0, requires_grad=True)
    dR, dt, ex = _SplitTheta.apply(theta)
    assert dR.shape == (1, 6) and dt.shape == (1, 3) and ex.shape == (1, 2)
    (2 * dR.sum() + 3 * dt.sum() + 5 * ex.sum()).backward()
    assert torch.equal(theta.grad, torch.tensor([2.0] * 6 + [3.0] * 3 + [5.0] * 2))
    theta.grad = None
    dR, dt, ex = _SplitTheta.apply(theta)
    dt.sum().backward()                           # unused outputs give zero blocks
    assert torch.equal(theta.grad, torch.tensor([0.0] * 6 + [1.0] * 3 + [0.0] * 2))


def test_trajectory_is_rigid_and_starts_at_identity():
    T = smooth_trajectory(50)
    assert torch.equal(T[0], torch.eye(4))
    R = T[:, :3, :3]
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand(50, 3, 3), atol=1e-6)
    assert (T[1:, :3, 3] - T[:-1, :3, 3]).norm(dim=1).max() < 0.02


def test_tracker_refuses_a_cpu_device():
    with pytest.raises(RuntimeError, match="CUDA"):
        Tracker(Camera(torch.eye(3), 8, 8), device="cpu")


def test_affine_inverse_matches_the_general_inverse():
    T = smooth_trajectory(40)[1:]
    assert torch.allclose(affine_inverse(T), torch.linalg.inv(T), atol=1e-6)
    assert torch.allclose(affine_inverse(T[3]) @ T[3], torch.eye(4), atol=1e-6)
    skew = T[5].clone()
    skew[:3, :3] = skew[:3, :3] * torch.tensor([1.1, 0.9, 1.3]) + 0.05       # not a rotation
    assert torch.allclose(affine_inverse(skew), torch.linalg.inv(skew), atol=1e-5)


def test_constant_motion_recurrence_stays_orthonormal():
    # 300 chained extrapolations: the non-orthonormality must grow slowly (linearly), not explode
    step = smooth_trajectory(3)[1]
    a, b = torch.eye(4), step.clone()
    for _ in range(300):
        a, b = b, constant_motion(a, b)
    R = b[:3, :3]
    assert (R @ R.t() - torch.eye(3)).abs().max() < 5e-3       # torch.linalg.inv drifts alike (1e-3)
