"""Backend -> Frontend map hand-off across a PROCESS boundary.

The reference runs tracking and mapping as two spawned processes (main.py:61-94) and ships the whole map
on every sync: `Backend.sync` puts `self.splats.no_grad_clone()` (seven fresh allocations + copies) plus
deep copies of the keyframes into an `mp.Queue` (gslam/backend.py:508-519) -- pickled by torch's CUDA-IPC
reductions, one `cudaIpcGetMemHandle` per storage on the way in and one `cudaIpcOpenMemHandle` per storage
on the way out -- and `Frontend.sync` deep-copies it again (gslam/frontend.py:253-269).  Every sync
allocates, exports and maps new device memory in both processes.

Here the two processes share ONE allocation, exported once at start-up:

    buf[2, W * capacity]   two slots of the flat parameter layout (mapping.FlatSplats blocks of `capacity`
                           rows: means | quats | scales | opacities | colors | log_uncertainties)
    hdr[4] int64           hdr[0] = version << 32 | slot << 31 | n_live, written by ONE 8-byte store that is
                           stream-ordered after the slot's copy; hdr[1] = last version the consumer finished
    claim (host, shared)   the version the consumer is copying right now (0: none)

`MapPublisher.publish` is one device-to-device copy into the slot the consumer is not reading plus the header
store -- no allocation, no export, no host sync on the mapping stream; `MapSubscriber.poll` is an 8-byte read,
`snapshot` one device-to-device copy into the tracker's fixed-capacity buffers (`Tracker.set_map`), so the
tracker's captured CUDA graphs keep replaying across syncs.  An inter-process CUDA event recorded after every
publish lets the consumer's stream wait for the copy without either host blocking.

Double buffering: version v lives in slot v % 2.  The publisher may overwrite slot v % 2 (with v + 2) only
when the consumer is not in the middle of copying version v: the consumer announces the version it is about
to copy in `claim` (host shared memory, written before the copy is enqueued) and confirms it in hdr[1]
(stream-ordered after the copy); a publish that would tear a snapshot is skipped and reported (False).
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from .map import GaussianSplattingData
from .mapping import FlatSplats, param_widths

_N_BITS = 31


def _pack(version: int, slot: int, n: int) -> int:
    return (version << 32) | (slot << _N_BITS) | n


def _unpack(word: int) -> Tuple[int, int, int]:
    return word >> 32, (word >> _N_BITS) & 1, word & ((1 << _N_BITS) - 1)


@dataclass
class MapHandle:
    """What travels to the consumer process ONCE (through a torch.multiprocessing queue or pipe: the tensors
    are shared by CUDA IPC / shared host memory, not copied)."""
    buf: torch.Tensor
    hdr: torch.Tensor
    claim: torch.Tensor
    event_handle: bytes
    capacity: int
    K: int
    device_index: int


class MapPublisher:
    """Backend side.  Keep this object alive for as long as the consumer uses the handle."""

    def __init__(self, capacity: int, device="cuda", K: int = 1):
        self.dev = torch.device(device)
        if self.dev.type != "cuda":
            raise RuntimeError("gslam_b200.handoff needs a CUDA device (the hand-off is device memory)")
        if self.dev.index is None:
            self.dev = torch.device("cuda", torch.cuda.current_device())
        self.capacity, self.K = int(capacity), int(K)
        self.widths = param_widths(self.K)
        W = sum(self.widths)
        self.buf = torch.zeros(2, W * self.capacity, dtype=torch.float32, device=self.dev)
        self.hdr = torch.zeros(4, dtype=torch.int64, device=self.dev)
        self.claim = torch.zeros(2, dtype=torch.int64).share_memory_()
        self.event = torch.cuda.Event(enable_timing=False, interprocess=True)
        self.version = 0
        self._ack_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self._side = torch.cuda.Stream(device=self.dev)
        self.skipped = 0

    def handle(self) -> MapHandle:
        with torch.cuda.device(self.dev):
            self.event.record()
        return MapHandle(self.buf, self.hdr, self.claim, self.event.ipc_handle(), self.capacity, self.K,
                         self.dev.index)

    def _consumer_ack(self) -> int:
        """hdr[1], read on a side stream so that the mapping stream is not waited for."""
        with torch.cuda.stream(self._side):
            self._ack_host.copy_(self.hdr[1:2], non_blocking=True)
        self._side.synchronize()
        return int(self._ack_host[0])

    @torch.no_grad()
    def publish(self, splats: FlatSplats) -> bool:
        """Copy the map into the free slot and flip the header.  False (nothing written) if the consumer is
        in the middle of copying the version that lives in that slot."""
        n = splats.n
        if n > self.capacity:
            raise RuntimeError(f"MapPublisher: the map has {n} splats, the shared buffers hold {self.capacity}")
        if splats.K != self.K:
            raise ValueError("MapPublisher: colour layout (K) differs from the channel's")
        v = self.version + 1
        slot = v & 1
        claimed = int(self.claim[0])
        if claimed != 0 and claimed == v - 2 and self._consumer_ack() < claimed:
            self.skipped += 1
            return False
        dst = FlatSplats.views(self.buf[slot], n, self.capacity, self.K)
        for d, p in zip(dst, splats.params):
            d.copy_(p.detach(), non_blocking=True)
        # ONE 8-byte store, stream-ordered after the copies (a fill kernel: the value travels in the launch
        # arguments, so back-to-back publishes cannot race on a host staging buffer)
        self.hdr[0].fill_(_pack(v, slot, n))
        self.event.record()
        self.version = v
        return True


class MapSubscriber:
    """Frontend side, built from the MapHandle received at start-up."""

    def __init__(self, handle: MapHandle, event: Optional[torch.cuda.Event] = None):
        """`event`: the publisher's own event object when both ends live in ONE process (an IPC event handle
        can only be opened by another process)."""
        self.dev = torch.device("cuda", handle.device_index)
        self.buf, self.hdr, self.claim = handle.buf, handle.hdr, handle.claim
        self.capacity, self.K = handle.capacity, handle.K
        self.event = event if event is not None else torch.cuda.Event.from_ipc_handle(self.dev, handle.event_handle)
        self._word = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.version = 0

    def poll(self) -> Optional[Tuple[int, int, int]]:
        """(version, slot, n_live) if a version newer than the last snapshot is published, else None.
        One 8-byte device read."""
        self._word.copy_(self.hdr[0:1], non_blocking=True)
        torch.cuda.current_stream(self.dev).synchronize()
        v, slot, n = _unpack(int(self._word[0]))
        return (v, slot, n) if v > self.version else None

    def view(self, slot: int, n: int) -> GaussianSplattingData:
        """The published map as zero-copy views of the shared slot (frozen: no gradients)."""
        P = FlatSplats.views(self.buf[slot], n, self.capacity, self.K)
        g = GaussianSplattingData(P[0], P[1], P[2], P[3], P[4], P[5],
                                  torch.zeros(n, dtype=torch.long, device=self.dev))
        for name in GaussianSplattingData._per_splat_params:
            getattr(g, name).requires_grad_(False)
        return g

    @torch.no_grad()
    def snapshot(self, tracker) -> Optional[int]:
        """If a new version is published: `tracker.set_map` from it (one device-to-device copy into the
        tracker's fixed-capacity buffers; its CUDA graphs survive) -> the version; else None."""
        got = self.poll()
        if got is None:
            return None
        v, slot, n = got
        self.claim[0] = v                                   # before the copy is enqueued
        stream = torch.cuda.current_stream(self.dev)
        stream.wait_event(self.event)                       # everything the publisher recorded so far
        tracker.set_map(self.view(slot, n))
        self.hdr[1].fill_(v)                                # stream-ordered after the copy: "finished v"
        self.version = v
        return v

    def wait_for(self, version: int, timeout_s: float = 30.0) -> Tuple[int, int, int]:
        """Block (polling) until `version` or a later one is published."""
        t0 = time.perf_counter()
        while True:
            self._word.copy_(self.hdr[0:1], non_blocking=True)
            torch.cuda.current_stream(self.dev).synchronize()
            v, slot, n = _unpack(int(self._word[0]))
            if v >= version:
                return v, slot, n
            if time.perf_counter() - t0 > timeout_s:
                raise TimeoutError(f"MapSubscriber: version {version} not published within {timeout_s} s")
            time.sleep(0.0005)
