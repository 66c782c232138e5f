"""ctypes binding of libgsb.so (the C ABI declared in include/gsb.h).

There is NO fallback: if the library is missing or a call fails this module
raises.  The oracle under oracle/ is never imported from here.
"""
from __future__ import annotations

import ctypes as ct
import os
import re
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSB_LIB overrides the path (kernel A/B experiments); the default is the in-tree build
LIB_PATH = os.environ.get("GSB_LIB") or os.path.join(_HERE, "libgsb.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "gsb.h")

_p = ct.c_void_p
_i = ct.c_int
_l = ct.c_int64
_f = ct.c_float
_z = ct.c_size_t

# name -> (restype, argtypes); mirrors include/gsb.h one to one
SIGNATURES = {
    "gsb_version": (_i, []),
    "gsb_last_error": (ct.c_char_p, []),
    "gsb_read_i64_host": (_i, [_p, _p, _p]),
    "gsb_project_fwd": (_i, [_p] * 8 + [_i, _i, _i, _i, _f, _f, _f, _f, _i, _i] + [_p] * 5 + [_p, _i, _i] + [_p]),
    "gsb_sh_bwd": (_i, [_p, _p, _i, _i, _p, _p, _p, _i, _i, _p, _p, _l, _p]),
    "gsb_tile_scan": (_i, [_p, _i, _i, _i, _p, _p, _p, _p]),
    "gsb_isect_scatter": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _l, _p, _p, _p]),
    "gsb_segment_sort": (_i, [_p, _p, _p, _i, _i, _i, _i, _l, _p, _p, _p]),
    "gsb_project_bwd": (_i, [_p] * 8 + [_i, _i, _i, _i, _f] + [_p, _p, _p, _l, _p, _i] + [_p] * 8 + [_l, _p] + [_p]),
    "gsb_project_bwd_det_floats": (_z, [_i, _i]),
    "gsb_pack_visible": (_i, [_p] * 4 + [_i, _i] + [_p] * 5 + [_p]),
    "gsb_scan_workspace_bytes": (_z, [_l]),
    "gsb_scan_i32_i64": (_i, [_p, _l, _i, _p, _p, _z, _p]),
    "gsb_isect_emit": (_i, [_p] * 4 + [_i, _i, _i, _i, _i, _p, _p, _p]),
    "gsb_sort_workspace_bytes": (_z, [_l]),
    "gsb_sort_pairs": (_i, [_p] * 4 + [_l, _i, _p, _z, _p]),
    "gsb_isect_offsets": (_i, [_p, _l, _i, _i, _i, _i, _p, _p]),
    "gsb_composite_fwd": (_i, [_p] * 4 + [_i] * 6 + [_l, _f] + [_p] * 5 + [_p]),
    "gsb_ssim_fwd": (_i, [_p] * 4 + [_i] * 5 + [_p] * 5 + [_p]),
    "gsb_ssim_fwd_det_floats": (_z, [_i, _i, _i, _i]),
    "gsb_ssim_bwd": (_i, [_p] * 4 + [_i] * 4 + [_p] * 4 + [_f, _p, _p, _i, _p]),
    "gsb_photo_loss": (_i, [_p, _p, _p, _i, _i, _i, _i, _f, _p, _p, _p, _p, _p, _p]),
    "gsb_photo_loss_det_floats": (_z, [_i, _i, _i]),
    "gsb_scale_by_scalar": (_i, [_p, _l, _p, _p]),
    "gsb_adam_multi": (_i, [_p, _i, _f, _f, _f, _i, _i, _p, _i, _f, _p, _p]),
    "gsb_composite_bwd": (_i, [_p] * 4 + [_i] * 6 + [_l] + [_p] * 6 + [_p, _p, _p, _z] + [_p]),
    "gsb_composite_bwd_det_bytes": (_z, [_l]),
    "gsb_map_update": (_i, [_p, _i, _p, _l, _i, _l, _l, _i, _f, _i, _f, _f, _f, _f, _i, _p, _p, _p, _p, _p]),
    "gsb_isotropic_reg_static_max_n": (_l, []),
    "gsb_isotropic_reg_static": (_i, [_p, _p, _l, _f, _p, _p, _p]),
    "gsb_map_update_mc_scratch_floats": (_z, [_l, _l]),
    "gsb_map_update_mc": (_i, [_p, _p, _i, _p, _l, _i, _l, _l, _i, _f, _i, _f, _f, _f, _f, _i, _p, _p, _p, _p, _i, _p, _p]),
    "gsb_vis_count": (_i, [_p, _i, _l, _p, _p]),
    "gsb_isotropic_reg": (_i, [_p, _p, _l, _f, _p, _p, _p, _p]),
    "gsb_rows_resize": (_i, [_p, _i, _l, _p, _l, _l, _p]),
    "gsb_lbfgs_state_bytes": (_z, []),
    "gsb_lbfgs_init": (_i, [_p, _p, _i, _i, _i, _i, ct.c_double, ct.c_double, ct.c_double, _p]),
    "gsb_lbfgs_advance": (_i, [_p, _p, _p, _p, _p]),
    "gsb_pose_fwd": (_i, [_p, _p, _p, _i, _p, _p]),
    "gsb_pose_bwd": (_i, [_p, _p, _p, _i, _p, _p, _p]),
}



class RowTensor(ct.Structure):
    """gsb_row_tensor of include/gsb.h."""
    _fields_ = [("src", _p), ("dst", _p), ("add", _p), ("width", _i)]


class MapTensor(ct.Structure):
    """gsb_map_tensor of include/gsb.h."""
    _fields_ = [("param", _p), ("exp_avg", _p), ("exp_avg_sq", _p), ("width", _i), ("col", _i), ("lr", _f)]


class AdamTensor(ct.Structure):
    """gsb_adam_tensor of include/gsb.h."""
    _fields_ = [("param", _p), ("grad", _p), ("exp_avg", _p), ("exp_avg_sq", _p), ("n", _l), ("lr", _f)]


_lib: Optional[ct.CDLL] = None


def header_symbols() -> list[str]:
    """Every function name declared in include/gsb.h."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsb_[a-z0-9_]+)\s*\(", src)))


def lib() -> ct.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make` (or __graft_entry__.build()). "
                "gslam_b200 has no CPU or PyTorch fallback.")
        L = ct.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class GsbError(RuntimeError):
    pass


def check(code: int, what: str) -> None:
    if code != 0:
        msg = lib().gsb_last_error().decode(errors="replace")
        raise GsbError(f"{what} failed with code {code}: {msg}")


# ---- launch accounting / per-call device timing (used by bench.py) -----------------------
launch_count = 0          # number of kernel-launching C-ABI calls made so far
_timing: Optional[dict] = None   # name -> list[(start_event, end_event)] while profiling

# device kernels each entry point enqueues (cub scan: 2, cub radix sort: see sort.cu)
KERNELS_PER_CALL = {"gsb_read_i64_host": 0}


def call(name: str, *args) -> None:
    """Invoke a kernel-launching entry point, count it, optionally time it with CUDA events."""
    global launch_count
    fn = getattr(lib(), name)
    launch_count += 1
    if _timing is None:
        check(fn(*args), name)
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(fn(*args), name)
    e1.record()
    _timing.setdefault(name, []).append((e0, e1))


def start_timing() -> None:
    global _timing
    _timing = {}


def stop_timing() -> dict:
    """-> {entry point: (calls, mean milliseconds per call)}; synchronises."""
    global _timing
    t, _timing = _timing or {}, None
    torch.cuda.synchronize()
    return {k: (len(v), sum(a.elapsed_time(b) for a, b in v) / len(v)) for k, v in t.items()}


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def ptr(t: Optional[torch.Tensor]):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_cuda, "gslam_b200 kernels need CUDA tensors (there is no CPU path)"
    # kernels are enqueued on the CURRENT device's current stream (stream() below): a tensor of another GPU would
    # be an illegal address there, or a silent wrong-GPU run with peer access -- fail loudly instead
    if _raw_device is not None and t.device.index != _raw_device():
        raise RuntimeError(f"gslam_b200: tensor on cuda:{t.device.index} but the current device is "
                           f"cuda:{_raw_device()}; wrap the call in torch.cuda.device(tensor.device)")
    return t.data_ptr()


def stream() -> int:
    """cudaStream_t of torch's current stream on the current device.  The raw accessors are ~20x
    cheaper than torch.cuda.current_stream() (17 us a call, 10 calls per step: a quarter of the host
    time of an eager step); the public path is the fallback if a torch build lacks them."""
    if _raw_stream is not None and _raw_device is not None:
        return _raw_stream(_raw_device())
    return torch.cuda.current_stream().cuda_stream


_workspaces: dict = {}


def workspace(kind: str, nbytes: int, device) -> torch.Tensor:
    """A cached scratch buffer per (kind, device, stream), grown geometrically."""
    key = (kind, str(device), stream())
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes * 1.5), 1 << 16), dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf
