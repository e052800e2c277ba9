#!/usr/bin/env python
"""Drive get_mesh (marching cubes) and the depth-to-mesh batch on the C2 room scene a few times: the workload the
ncu launch lists / captures of those kernels are taken on.

    python tools/run_mesh.py [--frames 24] [--reps 3] [--d2m-frames 120] [--what mc,d2m]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--frames", type=int, default=24)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--d2m-frames", type=int, default=120)
    ap.add_argument("--what", default="mc,d2m")
    args = ap.parse_args()
    what = args.what.split(",")
    frames = bench.make_frames(args.config, "room", args.frames)
    import torch
    from hive_b200 import geometric
    from hive_b200 import marching_cubes as mc
    from hive_b200.fusion import TSDFVolume
    sc = bench._scene(args.config, "room")
    if "mc" in what:
        vol = TSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
        dev = vol.upload_frames(frames, sc.K)
        for i in range(len(dev)):
            vol.integrate_device(dev, i)
        torch.cuda.synchronize()
        for r in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            mesh = mc.extract_device_mesh(vol)
            e1.record()
            torch.cuda.synchronize()
            print(f"mc rep {r}: {e0.elapsed_time(e1):.3f} ms, {mesh.verts.shape[0]} vertices, {mesh.faces.shape[0]} triangles",
                  flush=True)
    if "d2m" in what:
        B = args.d2m_frames
        fr = [sc.frame(i) for i in range(min(12, B))]
        depths = np.stack([fr[i % len(fr)].depth for i in range(B)])
        masks = np.stack([(fr[i % len(fr)].mask > 0) for i in range(B)]).astype(np.uint8)
        poses = np.stack([np.linalg.inv(fr[i % len(fr)].pose) for i in range(B)])
        dev = torch.device("cuda", torch.cuda.current_device())
        d, m = torch.from_numpy(depths).to(dev), torch.from_numpy(masks).to(dev)
        cams = torch.from_numpy(np.concatenate([poses[:, :3, :3].reshape(B, 9), poses[:, :3, 3].reshape(B, 3)], axis=1)).to(dev)
        kinv = geometric._kinv64(sc.K)
        for r in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = geometric.depth_to_mesh_device(d, m, kinv, cams, 0.1, 1)
            e1.record()
            torch.cuda.synchronize()
            print(f"d2m rep {r}: {e0.elapsed_time(e1):.3f} ms, {out.verts.shape[0]} vertices, {out.faces.shape[0]} faces", flush=True)


if __name__ == "__main__":
    main()
