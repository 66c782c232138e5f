"""Camera pose parameterisation of the hot path: the reference's `PoseZhou` and `Camera`.

Mirrors reference gslam/primitives.py:15-36 (`rotation_6d_to_matrix`), :40-100 (`PoseZhou`),
:350-367 (`Camera`, the three fields the rasterizer path reads).  `PoseZhou.forward()` is the
gradient sink of `v_viewmats`: the reference spends ~25 autograd nodes per view per iteration on it;
here the whole chain is one kernel forward (`gsb_pose_fwd`) and one backward (`gsb_pose_bwd`), and
`pose_batch` runs every view of a window in the same launch.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List

import torch
from torch import Tensor

from ._lib import call, ptr, stream


class _PoseChain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, Rt: Tensor, dR: Tensor, dt: Tensor) -> Tensor:
        Rt, dR, dt = Rt.contiguous().float(), dR.contiguous().float(), dt.contiguous().float()
        C = dR.shape[0]
        out = torch.empty(C, 4, 4, device=dR.device, dtype=torch.float32)
        call("gsb_pose_fwd", ptr(Rt), ptr(dR), ptr(dt), C, ptr(out), stream())
        ctx.save_for_backward(Rt, dR)
        return out

    @staticmethod
    def backward(ctx, v_viewmats: Tensor):
        Rt, dR = ctx.saved_tensors
        C = dR.shape[0]
        need_dR, need_dt = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        if not (need_dR or need_dt):
            return None, None, None
        v = v_viewmats.contiguous().float()
        v_dR = torch.empty_like(dR) if need_dR else None
        v_dt = torch.empty(C, 3, device=dR.device, dtype=torch.float32) if need_dt else None
        call("gsb_pose_bwd", ptr(Rt), ptr(dR), ptr(v), C, ptr(v_dR), ptr(v_dt), stream())
        return None, v_dR, v_dt


def pose_chain(Rt: Tensor, dR: Tensor, dt: Tensor) -> Tensor:
    """viewmats [C,4,4] = Rt [C,4,4] @ [[GS(dR [C,6] + id6), dt [C,3]], [0 0 0 1]]  (CUDA only)."""
    if not dR.is_cuda:
        raise RuntimeError("gslam_b200.primitives.pose_chain needs CUDA tensors (no CPU fallback)")
    return _PoseChain.apply(Rt.reshape(-1, 4, 4), dR.reshape(-1, 6), dt.reshape(-1, 3))


def rotation_6d_to_matrix(d6: Tensor) -> Tensor:
    """Zhou et al. 6-D rotation -> matrix, rows (b1, b2, b3); gslam/primitives.py:15-36."""
    lead = d6.shape[:-1]
    flat = d6.reshape(-1, 6)
    ident = torch.tensor([1.0, 0, 0, 0, 1.0, 0], device=d6.device, dtype=d6.dtype)
    eye = torch.eye(4, device=d6.device, dtype=torch.float32).expand(flat.shape[0], 4, 4)
    zero = torch.zeros(flat.shape[0], 3, device=d6.device, dtype=torch.float32)
    return pose_chain(eye, flat - ident, zero)[:, :3, :3].reshape(*lead, 3, 3)


class PoseZhou(torch.nn.Module):
    """Same attributes and behaviour as the reference class (gslam/primitives.py:40-92)."""

    def __init__(self, _pose: Tensor = None, is_learnable: bool = True):
        super().__init__()
        self.is_learnable = is_learnable
        self.Rt = torch.nn.Buffer(torch.eye(4) if _pose is None else _pose)
        self.dt = torch.nn.Parameter(torch.zeros(3, dtype=torch.float32), requires_grad=is_learnable)
        self.dR = torch.nn.Parameter(torch.zeros(6, dtype=torch.float32), requires_grad=is_learnable)
        self.register_buffer("id", torch.tensor([1, 0, 0, 0, 1, 0], device=self.Rt.device, dtype=torch.float32))
        self.register_buffer("eye4_row4", torch.tensor([0, 0, 0, 1], device=self.Rt.device, dtype=self.Rt.dtype))

    def forward(self) -> Tensor:
        if not self.is_learnable:
            return self.Rt
        return pose_chain(self.Rt, self.dR, self.dt)[0]


def pose_batch(poses: List[PoseZhou]) -> Tensor:
    """create_batch(poses, lambda p: p()) of gslam/map.py:86 with ONE pose kernel for the window."""
    if not all(isinstance(p, PoseZhou) and p.is_learnable for p in poses):
        return torch.stack([p() for p in poses], dim=0)
    return pose_chain(torch.stack([p.Rt for p in poses]), torch.stack([p.dR for p in poses]),
                      torch.stack([p.dt for p in poses]))


def affine_inverse(T: Tensor) -> Tensor:
    """Inverse of [A t; 0 0 0 1] (shape [...,4,4]) in closed form, [A^-1 | -A^-1 t] with A^-1 from the
    adjugate: what the reference gets from `torch.linalg.inv` on its view matrices
    (gslam/frontend.py:183, gslam/backend.py:764, gslam/insertion.py:218) without the LU
    factorisation, its pivoting sync and the ~170 ms solver start-up.  Deliberately NOT the
    transpose shortcut of a rigid matrix: the constant-motion recurrence b a^-1 b amplifies the
    rounding-level non-orthonormality of its inputs exponentially when a^T stands in for a^-1."""
    A = T[..., :3, :3]
    r0, r1, r2 = A[..., 0, :], A[..., 1, :], A[..., 2, :]
    c0, c1, c2 = torch.linalg.cross(r1, r2), torch.linalg.cross(r2, r0), torch.linalg.cross(r0, r1)
    det = (r0 * c0).sum(-1, keepdim=True)
    Ainv = torch.stack([c0, c1, c2], dim=-1) / det.unsqueeze(-1)
    out = torch.zeros_like(T)
    out[..., :3, :3] = Ainv
    out[..., :3, 3] = -(Ainv @ T[..., :3, 3:4])[..., 0]
    out[..., 3, 3] = 1.0
    return out


@dataclass
class Camera:
    """The fields of the reference Camera (gslam/primitives.py:350-367) that the render path reads."""
    intrinsics: Tensor
    height: int
    width: int

    def to(self, device):
        self.intrinsics = self.intrinsics.to(device)
        return self

    def clone(self):
        return Camera(self.intrinsics.detach(), self.height, self.width)
