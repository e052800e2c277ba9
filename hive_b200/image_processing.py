"""GPU drop-in for ``hive.image_processing.dilate_mask`` (hive/image_processing.py:32-48) and the batched helpers the
TSDF driver uses around it.  Host numpy in, host numpy out; no CPU fallback.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from . import _xfer

__all__ = ["MaskDilationOptions", "dilate_mask", "dilate_masks_device", "depth_max"]

_RECT3 = np.ones((3, 3), dtype=np.uint8)  # cv2.getStructuringElement(cv2.MORPH_RECT, (3, 3)), hive/options.py:248


class MaskDilationOptions:
    """hive/options.py:245-254: ``num_iterations`` applications of ``filter`` (default: the 3x3 rectangle)."""

    def __init__(self, num_iterations=0, dilation_filter=_RECT3):
        self.num_iterations = num_iterations
        self.filter = dilation_filter


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _check_filter(options) -> int:
    f = np.asarray(options.filter)
    if f.shape != (3, 3) or not np.all(f != 0):
        raise NotImplementedError("hive_b200.dilate_mask implements HIVE's default 3x3 rectangular filter only "
                                  "(hive/options.py:248)")
    n = int(options.num_iterations)
    assert n >= 0, "num_iterations must be >= 0"
    return n


def dilate_masks_device(masks, iterations: int):
    """(B, H, W) uint8 device tensor -> (B, H, W) uint8 0/1 device tensor, dilated by a (2*iterations+1)^2 square."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    B, H, W = (int(s) for s in masks.shape)
    dev = masks.device
    with torch.cuda.device(dev):
        nbytes = ctypes.c_int64(0)
        _lib.check(lib.hv_dilate_scratch_bytes(B, H, W, ctypes.byref(nbytes)))
        scratch = torch.empty(int(nbytes.value) // 4 + 1, dtype=torch.int32, device=dev)
        out = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
        _lib.check(lib.hv_dilate_mask(_ptr(masks), B, H, W, int(iterations), _ptr(scratch), int(nbytes.value), _ptr(out),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def dilate_mask(mask, dilation_options: MaskDilationOptions):
    """hive/image_processing.py:32-48: (H, W) mask of any dtype -> (H, W) bool, dilated."""
    torch = _lib.require_cuda()
    mask = np.asarray(mask)
    assert len(mask.shape) == 2, f"mask was expected to have 2 dimensions but got {len(mask.shape)}."
    n = _check_filter(dilation_options)
    dev = torch.device("cuda", torch.cuda.current_device())
    nz = mask if mask.dtype in (np.uint8, np.bool_) else (mask != 0)  # astype(float32) keeps "non-zero" for any dtype
    d = _xfer.to_device(nz[None], dev, np.uint8)
    return _xfer.to_host(dilate_masks_device(d, n))[0].view(np.bool_)


def depth_max(depths):
    """np.max per frame of a (B, H, W) float32 device tensor or host array (fusion.py:364), as a host (B,) array."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    if isinstance(depths, np.ndarray):
        depths = _xfer.to_device(depths, torch.device("cuda", torch.cuda.current_device()), np.float32)
    B = int(depths.shape[0])
    n = int(depths[0].numel())
    dev = depths.device
    with torch.cuda.device(dev):
        keys = torch.empty(B, dtype=torch.int32, device=dev)
        out = torch.empty(B, dtype=torch.float32, device=dev)
        _lib.check(lib.hv_depth_max(_ptr(depths), B, n, _ptr(keys), _ptr(out),
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return _xfer.to_host(out)
