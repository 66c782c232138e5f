"""CUDA-graph capture of a whole fwd+bwd step (launch-bound loops: tracking closure, mapping step).

The only host sync of the eager path is the intersection-count readback that
sizes the key buffers (upstream gsplat has the same sync inside isect_tiles).
Under capture the buffers get a fixed capacity instead and the kernels read the
count from device memory, so the whole step replays with one launch.  If a
replay ever produces more intersections than the capacity, `overflow` (a device
flag, sticky until `reset_overflow()`) is non-zero from that replay on and the caller must re-capture with a larger
capacity; results of an overflowed replay are incomplete.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from . import _lib, ops


_streams: dict = {}


def _capture_stream(device) -> torch.cuda.Stream:
    """One capture stream per device for the whole process (autograd's AccumulateGrad nodes remember
    the stream they first ran on; a fresh stream per capture would trip their stream check)."""
    key = torch.device(device).index
    if key not in _streams:
        _streams[key] = torch.cuda.Stream(device=device)
    return _streams[key]


_pools: dict = {}


def shared_pool(device):
    """A graph memory pool that outlives the graphs captured into it (a one-kernel keeper graph holds
    it open).  A sequence loop captures a new graph after every map change; with a private pool each
    capture cudaMallocs its whole working set again, with this one it takes the blocks the previous
    graph gave back.  Graphs that share a pool must never replay concurrently -- they all replay on
    the one stream of the loop."""
    key = torch.device(device).index
    if key not in _pools:
        handle = torch.cuda.graph_pool_handle()
        keeper = torch.cuda.CUDAGraph()
        cur = torch.cuda.current_stream(device)
        side = _capture_stream(device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            keeper.capture_begin(pool=handle)
            try:
                torch.zeros(1, device=device)
            finally:
                keeper.capture_end()
        cur.wait_stream(side)
        _pools[key] = (handle, keeper)
    return _pools[key][0]


class GraphedStep:
    def __init__(self, fn: Callable[[], torch.Tensor], headroom: float = 1.3, warmup: int = 2,
                 n_isects_hint: Optional[int] = None, pool=None):
        """`fn()` runs one full step (forward, loss, backward, collectives) on static tensors
        and returns the loss.  It is run eagerly `warmup` times (which also measures the
        intersection count), then captured.  A caller that has just run the same step eagerly
        passes the count it saw as `n_isects_hint` instead, and nothing runs before the capture.
        `pool`: a graph memory pool to capture into (see `shared_pool`); default a private one."""
        assert ops.isect_capacity is None
        n_max = int(n_isects_hint or 0)
        for _ in range(max(warmup, 1) if n_isects_hint is None else 0):
            fn()
            n_max = max(n_max, int(ops.last_n_isects.item()))
        self.capacity = int(n_max * headroom) + 4096
        self.overflow = torch.zeros(1, dtype=torch.int32, device=ops.last_n_isects.device)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        launches0 = _lib.launch_count
        ops.isect_capacity = self.capacity
        try:
            # capture_begin / capture_end on a side stream, i.e. `with torch.cuda.graph(...)` without its
            # gc.collect() + empty_cache(): a sequence re-captures the mapping step after every map
            # change, and emptying the allocator cache each time made every following allocation a
            # cudaMalloc (measured in bench.py's 60-frame sequence, ~20 captures: mapping time 1.48 s -> 0.67 s)
            cur = torch.cuda.current_stream()
            side = _capture_stream(cur.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                self.graph.capture_begin(pool=pool) if pool is not None else self.graph.capture_begin()
                try:
                    # a step that renders nothing (a rank without a view of a sharded window) must not pick up the
                    # flag of whatever rasterization() call came last -- a tensor of another graph, or freed memory
                    ops.last_overflow = None
                    self.loss = fn()
                    # every rasterization() call starts from a cleared flag, so the flag of a replay is
                    # gone with the next one: keep a sticky copy that only reset_overflow() clears
                    if ops.last_overflow is not None:
                        torch.maximum(self.overflow, ops.last_overflow, out=self.overflow)
                finally:
                    self.graph.capture_end()
            cur.wait_stream(side)
            self.n_isects = ops.last_n_isects
        finally:
            ops.isect_capacity = None
        self.launches_per_replay = _lib.launch_count - launches0

    def replay(self) -> torch.Tensor:
        self.graph.replay()
        return self.loss

    def reset_overflow(self) -> None:
        self.overflow.zero_()

    def check(self) -> None:
        if int(self.overflow.item()) != 0:
            raise RuntimeError(
                f"GraphedStep: {int(self.n_isects.item())} tile intersections exceed the captured "
                f"capacity {self.capacity}; re-capture with more headroom")
