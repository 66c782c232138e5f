"""Single-launch Adam for the splat parameters (K12/K13).

`FusedAdam` is a `torch.optim.Optimizer` whose per-parameter state has the keys
torch's Adam uses (`step`, `exp_avg`, `exp_avg_sq`, leading dimension N), so the
reference's optimizer-state surgery on insertion / pruning (gslam/insertion.py:51-62,
gslam/pruning.py:36-45) works on it unchanged.  `step()` updates every parameter of
every group with ONE kernel launch (the reference steps seven optimizers,
gslam/backend.py:559-563) and can fold in the opacity decay of gslam/backend.py:356-359.
"""
from __future__ import annotations

import ctypes as ct
from typing import Optional

import torch

from ._lib import AdamTensor, MapTensor, call, stream


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 capturable: bool = False):
        """capturable=True keeps ONE step counter for all parameters on the device (int32, like
        torch's capturable Adam keeps `step` on the device), so `step()` can be captured in a CUDA
        graph and replayed; every parameter is then updated by every call."""
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self.capturable = capturable
        self._dev_step = None        # int32[2] on the parameters' device: {steps done, ticket}

    def _state(self, p):
        st = self.state[p]
        if len(st) == 0:
            if self.capturable:
                if self._dev_step is None:
                    self._dev_step = torch.zeros(2, dtype=torch.int32, device=p.device)
                st["step"] = self._dev_step[0:1]
            else:
                st["step"] = torch.tensor(0.0)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @torch.no_grad()
    def step(self, closure=None, decay_param: Optional[torch.Tensor] = None,
             vis_count: Optional[torch.Tensor] = None, decay: float = 0.995):
        """decay_param (e.g. the opacity logits) is multiplied by `decay` where
        vis_count (int32, one entry per row of decay_param) > 1."""
        loss = closure() if closure is not None else None
        buckets = {}   # (step, betas, eps) -> list of (p, grad, state, lr)
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                    raise RuntimeError("FusedAdam needs contiguous float32 CUDA parameters")
                st = self._state(p)
                if self.capturable:
                    key = (0, tuple(group["betas"]), group["eps"])
                else:
                    st["step"] += 1
                    key = (int(st["step"]), tuple(group["betas"]), group["eps"])
                buckets.setdefault(key, []).append((p, p.grad.contiguous(), st, group["lr"]))
        if self.capturable and (len(buckets) > 1 or any(len(v) > 8 for v in buckets.values())):
            raise RuntimeError("FusedAdam(capturable=True) updates all parameters in one launch: "
                               "at most 8 tensors sharing betas and eps")
        for (step, betas, eps), items in buckets.items():
            for i in range(0, len(items), 8):
                chunk = items[i:i + 8]
                arr = (AdamTensor * len(chunk))()
                decay_idx, vis_div = -1, 1
                for k, (p, g, st, lr) in enumerate(chunk):
                    arr[k] = AdamTensor(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                                        st["exp_avg_sq"].data_ptr(), p.numel(), lr)
                    if decay_param is not None and p is decay_param:
                        decay_idx = k
                        vis_div = max(p.numel() // max(vis_count.numel(), 1), 1)
                vc = vis_count.data_ptr() if decay_idx >= 0 else None
                call("gsb_adam_multi", ct.cast(arr, ct.c_void_p), len(chunk), betas[0], betas[1], eps, step,
                     decay_idx, vc, vis_div, decay,
                     self._dev_step.data_ptr() if self.capturable else None, stream())
        return loss

    @torch.no_grad()
    def step_rows(self, params, rows: torch.Tensor, layout, n0: int, n1: int,
                  vis_out: Optional[torch.Tensor] = None, scales_index: int = -1, iso_weight: float = 0.0,
                  decay_index: int = -1, decay: float = 0.995) -> torch.Tensor:
        """The mapping step's fused update (gsb_map_update) of the Gaussians [n0, n1): `params` are the six
        splat parameters in `mapping.PARAM_NAMES` order, their gradient is `rows[n, layout.cols[k] + j]`
        (ops.RowLayout) and rows[n, layout.vis] the visibility count.  Adds the isotropic-regulariser
        gradient on params[scales_index], runs Adam with this optimizer's state and learning rates, decays
        params[decay_index] where the count is above one, writes the counts to vis_out (int32).  Returns
        the range's unweighted isotropic loss as a [1] tensor.  One launch; capturable like `step()`."""
        from . import ops
        lrs, hyper = {}, {}
        for group in self.param_groups:
            for p in group["params"]:
                lrs[id(p)] = group["lr"]
                hyper[id(p)] = (tuple(group["betas"]), group["eps"])
        dev = rows.device
        if params[0].shape[0] == 0:
            return torch.zeros(1, dtype=torch.float32, device=dev)
        if len({hyper.get(id(p)) for p in params}) != 1:
            raise RuntimeError("FusedAdam.step_rows: the splat parameters must share betas and eps")
        betas, eps = hyper[id(params[0])]
        arr = (MapTensor * len(params))()
        step = 0
        for k, p in enumerate(params):
            if id(p) not in lrs:
                raise RuntimeError("FusedAdam.step_rows: parameter is not in this optimizer")
            if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                raise RuntimeError("FusedAdam needs contiguous float32 CUDA parameters")
            st = self._state(p)
            if not self.capturable:
                st["step"] += 1
                step = int(st["step"])
            width = p.numel() // max(p.shape[0], 1) if p.shape[0] else layout.widths[k]
            arr[k] = MapTensor(p.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), width,
                               layout.cols[k], lrs[id(p)])
        iso = torch.zeros(1, dtype=torch.float32, device=dev)
        part = ops.det_scratch((max(n1 - n0, 1) + 255) // 256, dev) if ops.deterministic else None
        call("gsb_map_update", ct.cast(arr, ct.c_void_p), len(params), rows.data_ptr(), rows.shape[1], layout.vis,
             n0, n1, scales_index, float(iso_weight), decay_index, float(decay), betas[0], betas[1], eps, step,
             self._dev_step.data_ptr() if self.capturable else None,
             vis_out.data_ptr() if vis_out is not None else None, iso.data_ptr(),
             part.data_ptr() if part is not None else None, stream())
        return iso

    @torch.no_grad()
    def step_rows_mc(self, params, splats, rows: torch.Tensor, rows_mc: int, layout, n0: int, n1: int, vis_mc: int,
                     stats_in: torch.Tensor, stats_mc: int, rank: int, scratch: torch.Tensor,
                     scales_index: int = -1, iso_weight: float = 0.0, decay_index: int = -1,
                     decay: float = 0.995) -> None:
        """`step_rows` as ONE kernel that is also the reduce-scatter and the all-gather (gsb_map_update_mc):
        rows_mc / vis_mc / stats_mc are NVLink-multicast addresses of symmetric allocations, `splats.flat` is
        a symmetric allocation too (splats.symm.multicast_ptr).  The caller provides the barriers."""
        import ctypes
        lrs, hyper = {}, {}
        for group in self.param_groups:
            for p in group["params"]:
                lrs[id(p)] = group["lr"]
                hyper[id(p)] = (tuple(group["betas"]), group["eps"])
        betas, eps = hyper[id(params[0])]
        arr = (MapTensor * len(params))()
        mc = (ctypes.c_uint64 * len(params))()
        step = 0
        flat_mc = int(splats.symm.multicast_ptr)
        base = splats.flat.data_ptr()
        for k, p in enumerate(params):
            st = self._state(p)
            if not self.capturable:
                st["step"] += 1
                step = int(st["step"])
            arr[k] = MapTensor(p.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), layout.widths[k],
                               layout.cols[k], lrs[id(p)])
            mc[k] = flat_mc + (p.data_ptr() - base)
        call("gsb_map_update_mc", ct.cast(arr, ct.c_void_p), ct.cast(mc, ct.c_void_p), len(params), rows_mc,
             rows.shape[1], layout.vis, n0, n1, scales_index, float(iso_weight), decay_index, float(decay),
             betas[0], betas[1], eps, step, self._dev_step.data_ptr() if self.capturable else None,
             vis_mc, stats_in.data_ptr(), stats_mc, rank, scratch.data_ptr(), stream())
