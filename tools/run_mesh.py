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
            mesh = mc.extract_device_mesh(vol, timing_event=True)
            e1.record()
            torch.cuda.synchronize()
            done = vol._mc_plan.done_event
            dev_ms = f"{e0.elapsed_time(done):.3f}" if done is not None else "n/a"
            print(f"mc rep {r}: {e0.elapsed_time(e1):.3f} ms (kernels only {dev_ms} ms), {mesh.verts.shape[0]} vertices, "
                  f"{mesh.faces.shape[0]} triangles", flush=True)
        # phase split: classification (sign + count + scans) | host read of the totals | emit + vertices
        plan = mc.McPlan(vol._tsdf_vol_gpu, vol._color_vol_gpu, tuple(int(d) for d in vol._local_dim), vol._pitch)
        for r in range(args.reps):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ev[0].record()
            plan.count()
            ev[1].record()
            t1 = time.perf_counter()
            nv, nt, _, _ = plan.read_totals()
            t2 = time.perf_counter()
            ev[2].record()
            plan.emit(nv, nt, vol._voxel_size, vol._vol_origin)
            ev[3].record()
            torch.cuda.synchronize()
            t3 = time.perf_counter()
            print(f"mc phases rep {r}: count {ev[0].elapsed_time(ev[1]):.3f} ms | totals gap {ev[1].elapsed_time(ev[2]):.3f} ms"
                  f" | emit {ev[2].elapsed_time(ev[3]):.3f} ms || host: launch {1e3 * (t1 - t0):.3f}, wait {1e3 * (t2 - t1):.3f},"
                  f" emit+sync {1e3 * (t3 - t2):.3f}", flush=True)
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
