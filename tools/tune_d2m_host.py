#!/usr/bin/env python
"""Where the host-to-host depth-to-mesh pipeline spends its time: the raw legs (staging copy into pinned memory by n
threads, H2D, D2H) and the whole call for a few (chunk_frames, copy_threads).  120 frames of 640x480."""
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def med(f, n=5):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        f()
        ts.append(1e3 * (time.perf_counter() - t0))
    return float(np.median(ts[1:]))


def main():
    import torch
    from hive_b200 import geometric
    B = 120
    sc = bench._scene("C2", "room")
    fr = [sc.frame(i) for i in range(12)]
    depths = np.stack([fr[i % 12].depth for i in range(B)])
    masks = np.stack([(fr[i % 12].mask > 0) for i in range(B)]).astype(np.uint8)
    poses = np.stack([np.linalg.inv(fr[i % 12].pose) for i in range(B)])
    pin = torch.empty(depths.shape, dtype=torch.float32, pin_memory=True)
    dev = torch.empty(depths.shape, dtype=torch.float32, device="cuda")
    out = torch.empty(339_000_000 // 4, dtype=torch.float32, device="cuda")
    pin_out = torch.empty(out.shape, dtype=torch.float32, pin_memory=True)
    for n in (1, 2, 4, 8):
        with ThreadPoolExecutor(n) as pool:
            parts = np.array_split(np.arange(B), n)

            def stage():
                list(pool.map(lambda idx: np.copyto(pin.numpy()[idx[0]:idx[-1] + 1], depths[idx[0]:idx[-1] + 1]), parts))
            print(f"staging {depths.nbytes / 1e6:.0f} MB with {n} threads: {med(stage):.2f} ms")

    def h2d():
        dev.copy_(pin, non_blocking=True)
        torch.cuda.synchronize()

    def d2h():
        pin_out.copy_(out, non_blocking=True)
        torch.cuda.synchronize()

    def both():
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        with torch.cuda.stream(s1):
            dev.copy_(pin, non_blocking=True)
        with torch.cuda.stream(s2):
            pin_out.copy_(out, non_blocking=True)
        torch.cuda.synchronize()
    print(f"H2D {depths.nbytes / 1e6:.0f} MB: {med(h2d):.2f} ms;  D2H {out.numel() * 4 / 1e6:.0f} MB: {med(d2h):.2f} ms;  both at once: {med(both):.2f} ms")
    for _ in range(3):
        geometric._HOST_TRACE = []
        geometric._TRACE = []
        t0 = time.perf_counter()
        geometric.depth_to_mesh_host(depths, masks, sc.K, poses)
        t1 = time.perf_counter()
    tr, geometric._HOST_TRACE = geometric._HOST_TRACE, None
    tr = sorted(tr + [("-", "  device call: " + lab, t) for lab, t in geometric._TRACE], key=lambda r: r[2])
    geometric._TRACE = None
    print(f"one call {1e3 * (t1 - t0):.2f} ms; main-thread steps (ms since the call began):")
    for c, step, t in tr:
        print(f"  chunk {c} {step:22s} {1e3 * (t - t0):7.2f}")
    for chunk in (20,):
        for threads in (4, 8, 12):
            ms = med(lambda: geometric.depth_to_mesh_host(depths, masks, sc.K, poses, chunk_frames=chunk, copy_threads=threads), 6)
            print(f"depth_to_mesh_host chunk {chunk:3d} threads {threads}: {ms:.2f} ms = {B / ms * 1e3:.0f} frames/s")


if __name__ == "__main__":
    main()
