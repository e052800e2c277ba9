"""Host <-> device transfers through pinned memory.

The reference-facing calls take and return host numpy arrays; a pageable ``.cpu()`` / ``.to(dev)``
stages through the driver's small bounce buffers and faults in fresh pages on every call, which costs several times the
kernels it surrounds (get_mesh at 512^3: 0.7 ms of kernels, 18 ms of pageable copies).  Here every transfer goes through
a page-locked block of torch's caching host allocator: the first call of a given size pays ``cudaHostAlloc``, later
calls reuse the block as soon as the numpy array that was handed out has been garbage collected, so results are always
fresh arrays (never aliased while alive) and copies run at PCIe/NVLink-C2C speed.
"""
from __future__ import annotations

import numpy as np

from . import _lib

_NP2TORCH = None


def _torch_dtype(torch, np_dtype):
    global _NP2TORCH
    if _NP2TORCH is None:
        _NP2TORCH = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
                     np.dtype(np.int32): torch.int32, np.dtype(np.int64): torch.int64,
                     np.dtype(np.uint8): torch.uint8, np.dtype(np.bool_): torch.uint8,
                     np.dtype(np.int16): torch.int16, np.dtype(np.uint16): torch.int16}
    return _NP2TORCH[np.dtype(np_dtype)]


def to_host(*tensors):
    """Device tensors -> fresh numpy arrays (pinned), all copies in flight together, one stream synchronise.

    ``None`` entries pass through.  Returns a tuple (or the single array when called with one tensor).
    """
    import torch
    staged, dev = [], None
    for t in tensors:
        if t is None or not t.is_cuda:  # host tensors (the gloo tests of the stitching logic) pass straight through
            staged.append(t)
            continue
        dev = t.device
        if t.numel() == 0:
            staged.append(torch.empty(t.shape, dtype=t.dtype))
            continue
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        with torch.cuda.device(dev):
            h.copy_(t if t.is_contiguous() else t.contiguous(), non_blocking=True)
        staged.append(h)
    if dev is not None:
        with torch.cuda.device(dev):
            torch.cuda.current_stream().synchronize()
    out = tuple(None if h is None else h.numpy() for h in staged)
    return out[0] if len(out) == 1 else out


def to_device(array, device, dtype=None):
    """Host array -> device tensor: one pass converts/copies into a pinned block, then an async H2D on the current
    stream (the caching host allocator keeps the block alive until that copy has run)."""
    torch = _lib.require_cuda()
    array = np.asarray(array)
    dtype = np.dtype(array.dtype if dtype is None else dtype)
    if dtype == np.bool_:
        dtype = np.dtype(np.uint8)
    tdt = _torch_dtype(torch, dtype)
    if array.size == 0:
        return torch.empty(array.shape, dtype=tdt, device=device)
    h = torch.empty(array.shape, dtype=tdt, pin_memory=True)
    np.copyto(h.numpy().view(dtype), array, casting="unsafe")
    with torch.cuda.device(device):
        return h.to(device, non_blocking=True)
