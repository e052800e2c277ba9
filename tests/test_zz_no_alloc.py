"""hv_dilate_mask_static: the entry point that allocates nothing (its scratch is a fixed array inside the library) against
hv_dilate_mask with caller scratch.  Both are bit work: the outputs must be equal.  Collected last on purpose
(the name)."""
import ctypes

import numpy as np
import pytest

from hive_b200 import _lib

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _dilate_with_scratch(lib, torch, masks, n):
    B, H, W = (int(s) for s in masks.shape)
    nbytes = ctypes.c_int64(0)
    _lib.check(lib.hv_dilate_scratch_bytes(B, H, W, ctypes.byref(nbytes)))
    scratch = torch.empty(int(nbytes.value) // 4 + 1, dtype=torch.int32, device=masks.device)
    out = torch.empty_like(masks)
    _lib.check(lib.hv_dilate_mask(_ptr(masks), B, H, W, n, _ptr(scratch), int(nbytes.value), _ptr(out),
                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def _numpy_dilate(mask, n):
    """out = any(mask != 0) over the (2n+1)^2 window, border ignored (hive/image_processing.py:32-48)."""
    on = mask != 0
    H, W = on.shape
    pad = np.zeros((H + 2 * n, W + 2 * n), dtype=bool)
    pad[n:n + H, n:n + W] = on
    out = np.zeros_like(on)
    for dv in range(2 * n + 1):
        for du in range(2 * n + 1):
            out |= pad[dv:dv + H, du:du + W]
    return out.astype(np.uint8)


def test_static_scratch_dilation_equals_caller_scratch_twice_in_a_row():
    torch = _lib.require_cuda()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    limit = ctypes.c_int64(0)
    _lib.check(lib.hv_dilate_static_max_bytes(ctypes.byref(limit)))
    assert limit.value == 2 * 1024 * 1024
    rng = np.random.default_rng(20261105)
    # two calls in a row of different shapes: the second reuses the words the first left behind (ragged widths: the
    # last word of a row is partial; the second call's planes are smaller than the first's)
    cases = [((3, 97, 211), 3, 0.002), ((2, 60, 75), 1, 0.01), ((1, 480, 640), 9, 0.0005), ((1, 5, 33), 0, 0.3)]
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for (B, H, W), n, density in cases:
            host = (rng.random((B, H, W)) < density).astype(np.uint8) * rng.integers(1, 256, (B, H, W), dtype=np.uint8)
            masks = torch.from_numpy(host).to(dev)
            want = _dilate_with_scratch(lib, torch, masks, n)
            got = torch.full_like(masks, 7)
            before = torch.cuda.memory_allocated(dev)
            _lib.check(lib.hv_dilate_mask_static(_ptr(masks), B, H, W, n, _ptr(got), stream))
            assert torch.cuda.memory_allocated(dev) == before
            torch.cuda.synchronize(dev)
            assert torch.equal(got, want), f"static-scratch dilation differs for {(B, H, W)} n={n}"
            assert np.array_equal(got[0].cpu().numpy(), _numpy_dilate(host[0], n))
        # the same input again right after a different one: no state survives between calls
        (B, H, W), n, _ = cases[0]
        again = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
        host = (rng.random((B, H, W)) < 0.004).astype(np.uint8)
        masks = torch.from_numpy(host).to(dev)
        for _ in range(2):
            again.fill_(9)
            _lib.check(lib.hv_dilate_mask_static(_ptr(masks), B, H, W, n, _ptr(again), stream))
            assert torch.equal(again, _dilate_with_scratch(lib, torch, masks, n))


def test_static_scratch_rejects_sizes_beyond_the_array():
    torch = _lib.require_cuda()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        B, H, W = 5, 1080, 1920  # 2 * 5 * 1080 * 60 words = 2.47 MiB > 2 MiB
        masks = torch.zeros((B, H, W), dtype=torch.uint8, device=dev)
        out = torch.full_like(masks, 3)
        rc = lib.hv_dilate_mask_static(_ptr(masks), B, H, W, 1, _ptr(out), None)
        assert rc == _lib.HV_E_CAPACITY
        assert "built-in array" in _lib.last_error()
        with pytest.raises(_lib.HiveB200Error):
            _lib.check(rc)
        torch.cuda.synchronize(dev)
        assert int(out.min()) == 3 and int(out.max()) == 3  # nothing was launched
        # four such masks fit exactly under the limit
        _lib.check(lib.hv_dilate_mask_static(_ptr(masks), 4, H, W, 1, _ptr(out), None))
        torch.cuda.synchronize(dev)
        assert int(out[:4].max()) == 0 and int(out[4].min()) == 3
