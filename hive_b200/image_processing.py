"""GPU drop-in for ``hive.image_processing.dilate_mask`` (hive/image_processing.py:32-48) and the batched helpers the
TSDF driver uses around it.  Host numpy in, host numpy out; no CPU fallback.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from . import _xfer

__all__ = ["MaskDilationOptions", "dilate_mask", "dilate_masks_device", "depth_max", "srgb_vertex_color_lut",
           "apply_byte_lut_device", "vertex_colors_to_srgb"]

_RECT3 = np.ones((3, 3), dtype=np.uint8)  # cv2.getStructuringElement(cv2.MORPH_RECT, (3, 3)), hive/options.py:248


class MaskDilationOptions:
    """hive/options.py:245-254: ``num_iterations`` applications of ``filter`` (default: the 3x3 rectangle)."""

    def __init__(self, num_iterations=0, dilation_filter=_RECT3):
        self.num_iterations = num_iterations
        self.filter = dilation_filter


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _check_filter(options) -> int:
    """The dilation radius: ``num_iterations`` applications of an all-ones k x k rectangle (k odd) are one dilation by a
    square of radius ``num_iterations * (k // 2)`` -- OpenCV folds the iterations of a rectangular element into its size the
    same way.  HIVE only ever builds the 3 x 3 rectangle (hive/options.py:248, ``from_args`` passes no filter)."""
    f = np.asarray(options.filter)
    if f.ndim != 2 or f.shape[0] != f.shape[1] or f.shape[0] % 2 != 1 or not np.all(f != 0):
        raise NotImplementedError("hive_b200.dilate_mask implements all-ones square filters of odd size (HIVE's default is "
                                  "the 3x3 rectangle, hive/options.py:248); other structuring elements are not supported")
    n = int(options.num_iterations)
    assert n >= 0, "num_iterations must be >= 0"
    return n * (int(f.shape[0]) // 2)


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


def srgb_vertex_color_lut():
    """The colour-space conversion of hive/pipeline.py:281-282 tabulated over its whole domain (one uint8 in, one out):
    ``(255 * np.power(vertex_colors / 255, 2.2)).astype(int)`` stored back into the uint8 colour array."""
    vertex_colors = np.arange(256, dtype=np.uint8)
    return (255 * np.power(vertex_colors / 255, 2.2)).astype(int).astype(np.uint8)


_LUT_CACHE = {}


def apply_byte_lut_device(bytes_tensor, lut):
    """In place ``t[...] = lut[t[...]]`` on a contiguous uint8 device tensor (``hv_byte_lut``)."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    assert bytes_tensor.dtype == torch.uint8 and bytes_tensor.is_contiguous(), "contiguous uint8 tensor expected"
    lut = np.ascontiguousarray(lut, dtype=np.uint8)
    assert lut.shape == (256,), "lut must have 256 entries"
    dev = bytes_tensor.device
    key = (dev.index, lut.tobytes())
    d_lut = _LUT_CACHE.get(key)
    if d_lut is None:
        d_lut = _LUT_CACHE[key] = _xfer.to_device(lut, dev, np.uint8)
    with torch.cuda.device(dev):
        _lib.check(lib.hv_byte_lut(_ptr(bytes_tensor), int(bytes_tensor.numel()), _ptr(d_lut),
                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return bytes_tensor


def vertex_colors_to_srgb(vertex_colors):
    """hive/pipeline.py:281-282 for a host (V, 3|4) uint8 colour array: returns the converted r, g, b (alpha, if any, is
    left alone as in the reference, which only assigns ``[:, :3]``)."""
    torch = _lib.require_cuda()
    vertex_colors = np.asarray(vertex_colors)
    assert vertex_colors.dtype == np.uint8 and vertex_colors.ndim == 2, "(V, 3) or (V, 4) uint8 colours expected"
    dev = torch.device("cuda", torch.cuda.current_device())
    rgb = _xfer.to_device(np.ascontiguousarray(vertex_colors[:, :3]), dev, np.uint8)
    out = vertex_colors.copy()
    out[:, :3] = _xfer.to_host(apply_byte_lut_device(rgb, srgb_vertex_color_lut()))
    return out
