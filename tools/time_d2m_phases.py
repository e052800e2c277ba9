#!/usr/bin/env python
"""Host-side phases of one batched depth-to-mesh call (where the time between the kernels goes): wall clock around each
statement of hive_b200.geometric.depth_to_mesh_device, with and without device syncs.  120 frames of 640x480."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    from hive_b200 import geometric
    B = 120
    sc = bench._scene("C2", "room")
    fr = [sc.frame(i) for i in range(12)]
    depths = np.stack([fr[i % 12].depth for i in range(B)])
    masks = np.stack([(fr[i % 12].mask > 0) for i in range(B)]).astype(np.uint8)
    poses = np.stack([np.linalg.inv(fr[i % 12].pose) for i in range(B)])
    dev = torch.device("cuda", 0)
    d, m = torch.from_numpy(depths).to(dev), torch.from_numpy(masks).to(dev)
    cams = torch.from_numpy(np.concatenate([poses[:, :3, :3].reshape(B, 9), poses[:, :3, 3].reshape(B, 3)], axis=1)).to(dev)
    kinv = geometric._kinv64(sc.K)
    for label in ("exactly sized (totals read between count and emit)", "capped (one read at the end)"):
        ev, wall = [], []
        for r in range(8):
            if label.startswith("exact"):
                geometric._D2M_CAPS.clear()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            geometric.depth_to_mesh_device(d, m, kinv, cams, 0.1, 1)
            e1.record()
            t1 = time.perf_counter()
            torch.cuda.synchronize()
            ev.append(e0.elapsed_time(e1))
            wall.append(1e3 * (t1 - t0))
        print(f"{label}: events median {np.median(ev[2:]):.3f} ms, wall median {np.median(wall[2:]):.3f} ms")
    steps = {}
    for r in range(6):
        torch.cuda.synchronize()
        geometric._TRACE = []
        geometric.depth_to_mesh_device(d, m, kinv, cams, 0.1, 1)
        tr = geometric._TRACE
        for a, b in zip(tr[:-1], tr[1:]):
            steps.setdefault(b[0], []).append(1e6 * (b[1] - a[1]))
    geometric._TRACE = None
    print("host steps, us (median): " + ", ".join(f"{k} {np.median(v[1:]):.0f}" for k, v in steps.items()))


if __name__ == "__main__":
    main()
