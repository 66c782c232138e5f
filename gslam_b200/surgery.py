"""Densify / prune with optimizer-state surgery on the device.

Mirrors reference gslam/pruning.py:10-53 (`prune_using_mask`) and gslam/insertion.py:27-62
(`_add_new_splats`) for the flat layout of `mapping.FlatSplats` + `optim.FusedAdam`: the six
parameters stay views of ONE [15 N] buffer, and so do `exp_avg` / `exp_avg_sq`, so the single-launch
Adam, the one-buffer gradient all-reduce and `project_bwd`'s flat gradient keep working after the map
has grown or shrunk.  Where the reference loops over 6 parameters x (param, exp_avg, exp_avg_sq) with
a boolean-index or `torch.cat` kernel each (plus the allocator traffic), this is one scan of the mask,
one 8-byte read of the new size, and ONE `gsb_rows_resize` launch over all 18 tensors.
"""
from __future__ import annotations

import ctypes as ct
from typing import Dict, List, Optional, Sequence

import torch
from torch import Tensor

from . import ops
from ._lib import RowTensor, call, ptr, stream
from .mapping import PARAM_NAMES, FlatSplats, _round_up



def _state_tensors(optimizer, params: Sequence[Tensor], key: str) -> List[Optional[Tensor]]:
    return [optimizer.state[p][key] if (optimizer is not None and p in optimizer.state
                                        and key in optimizer.state[p]) else None for p in params]


@torch.no_grad()
def resize_splats(splats: FlatSplats, optimizer, keep_mask: Optional[Tensor] = None,
                  new_params: Optional[Dict[str, Tensor]] = None,
                  per_gaussian_params: Optional[list] = None) -> int:
    """Drop the rows where keep_mask is False, then append new_params' rows.  Returns the new N.

    `optimizer` is a FusedAdam (or any optimizer with one parameter per group and the torch-Adam
    state keys); its state is re-keyed to the new parameter tensors, `step` is kept, the appended rows
    start with zero moments (gslam/insertion.py:55-59).  per_gaussian_params: extra [N, ...] tensors
    compacted with the same mask (gslam/pruning.py:48-51); they get plain torch indexing."""
    old_params = list(splats.params)
    n_old = splats.n
    dev = splats.flat.device
    W = splats.widths
    cum, n_keep = None, n_old
    if keep_mask is not None:
        assert keep_mask.shape == (n_old,)
        cum = ops.inclusive_scan(keep_mask.to(torch.int32).contiguous(), flag_positive=True)
        n_keep = ops.read_i64(cum) if n_old else 0       # the one host sync: sizes the new buffers
    n_add = 0
    adds: List[Optional[Tensor]] = [None] * 6
    if new_params is not None:
        n_add = new_params["means"].shape[0]
        for k, name in enumerate(PARAM_NAMES):
            t = new_params[name].detach().to(dev, torch.float32).contiguous()
            assert t.shape[0] == n_add and t.numel() == n_add * W[k], (name, tuple(t.shape))
            adds[k] = t
    n_new = n_keep + n_add

    # same block layout as FlatSplats: six blocks of `cap` rows (n rounded up to the rank count), padding zero
    cap_new = _round_up(n_new, splats.align)
    new_flat = torch.zeros(sum(W) * cap_new, device=dev, dtype=torch.float32)
    new_views = FlatSplats.views(new_flat, n_new, cap_new, splats.K)
    exp_avg, exp_avg_sq = _state_tensors(optimizer, old_params, "exp_avg"), _state_tensors(optimizer, old_params, "exp_avg_sq")
    has_state = any(t is not None for t in exp_avg)
    m_views = FlatSplats.views(torch.zeros(sum(W) * cap_new, device=dev), n_new, cap_new, splats.K) if has_state else [None] * 6
    v_views = FlatSplats.views(torch.zeros(sum(W) * cap_new, device=dev), n_new, cap_new, splats.K) if has_state else [None] * 6

    rows = []
    for k in range(6):
        rows.append((old_params[k].detach(), new_views[k], adds[k], W[k]))
        for src, dst in ((exp_avg[k], m_views[k]), (exp_avg_sq[k], v_views[k])):
            if src is not None:
                rows.append((src.contiguous(), dst, None, W[k]))
    arr = (RowTensor * len(rows))()
    for i, (src, dst, add, w) in enumerate(rows):
        arr[i] = RowTensor(ptr(src) if n_old else None, ptr(dst) if n_new else None, ptr(add), w)
    if n_new:
        call("gsb_rows_resize", ct.cast(arr, ct.c_void_p), len(rows), n_old, ptr(cum), n_keep, n_add, stream())

    # re-key: new leaf parameters (views of the new flat buffer), same state dict objects
    splats.flat, splats.n, splats.cap = new_flat, n_new, cap_new
    splats.params = [v.requires_grad_(True) for v in new_views]
    if optimizer is not None:
        by_old = {id(p): k for k, p in enumerate(old_params)}
        for group in optimizer.param_groups:
            group["params"] = [splats.params[by_old[id(p)]] if id(p) in by_old else p for p in group["params"]]
        for k, p_old in enumerate(old_params):
            if p_old in optimizer.state:
                st = optimizer.state.pop(p_old)
                if "exp_avg" in st:
                    st["exp_avg"], st["exp_avg_sq"] = m_views[k], v_views[k]
                optimizer.state[splats.params[k]] = st
    if per_gaussian_params is not None and keep_mask is not None:
        for i, t in enumerate(per_gaussian_params):
            per_gaussian_params[i] = t[keep_mask]
    return n_new


def prune_using_mask(splats: FlatSplats, optimizer, keep_mask: Tensor,
                     per_gaussian_params: Optional[list] = None) -> int:
    """gslam/pruning.py:10-53: returns the number of pruned splats; a mask that keeps nothing is a no-op."""
    n_old = splats.n
    if n_old == 0:
        return 0
    n_new = resize_splats(splats, optimizer, keep_mask=keep_mask, per_gaussian_params=per_gaussian_params) \
        if bool(keep_mask.any()) else n_old
    return n_old - n_new


def add_new_splats(splats: FlatSplats, optimizer, new_params: Dict[str, Tensor]) -> int:
    """gslam/insertion.py:27-62: append the rows of new_params (keys of PARAM_NAMES); returns how many."""
    n_add = new_params["means"].shape[0]
    if n_add:
        resize_splats(splats, optimizer, new_params=new_params)
    return n_add
