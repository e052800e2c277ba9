#!/usr/bin/env python
"""The reference's ``third_party/tsdf_fusion_python/demo.py`` with ``fusion`` replaced by ``hive_b200.fusion`` -- the
only line a user changes -- on seeded synthetic frames (the 7-Scenes frames of the reference do not travel).

    python tools/demo.py [--config C1] [--frames 100] [--voxel-size 0.02] [--out /tmp]

Same flow, same prints: frustum bounds over all frames, ``TSDFVolume(vol_bnds, voxel_size)``, ``integrate`` per frame with
the average FPS, ``get_mesh`` -> ``meshwrite('mesh.ply')``, ``get_point_cloud`` -> ``pcwrite('pc.ply')``.
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from hive_b200 import fusion  # noqa: E402   (the reference: `import fusion`)
from hive_b200.synthetic import Scene  # noqa: E402


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C1")
    ap.add_argument("--variant", default="room")
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--voxel-size", type=float, default=0.02)
    ap.add_argument("--out", default="/tmp")
    args = ap.parse_args()
    scene = Scene(args.config, args.variant, frames=args.frames)
    n_imgs = args.frames
    frames = [scene.frame(i) for i in range(n_imgs)]  # stands in for the cv2.imread / np.loadtxt calls of the demo

    print("Estimating voxel volume bounds...")
    cam_intr = scene.K
    vol_bnds = np.zeros((3, 2))
    for fr in frames:
        view_frust_pts = fusion.get_view_frustum(fr.depth, cam_intr, fr.pose)
        vol_bnds[:, 0] = np.minimum(vol_bnds[:, 0], np.amin(view_frust_pts, axis=1))
        vol_bnds[:, 1] = np.maximum(vol_bnds[:, 1], np.amax(view_frust_pts, axis=1))

    print("Initializing voxel volume...")
    tsdf_vol = fusion.TSDFVolume(vol_bnds, voxel_size=args.voxel_size)

    t0_elapse = time.time()
    for i, fr in enumerate(frames):
        if i % 25 == 0:
            print("Fusing frame %d/%d" % (i + 1, n_imgs))
        tsdf_vol.integrate(fr.rgb, fr.depth, cam_intr, fr.pose, obs_weight=1.)
    tsdf_vol.get_volume()  # the demo's FPS is wall clock: wait for the device before stopping it
    fps = n_imgs / (time.time() - t0_elapse)
    print("Average FPS: {:.2f}".format(fps))
    print("Volume: %s voxels" % "x".join(str(int(d)) for d in tsdf_vol._vol_dim))

    print("Saving mesh to mesh.ply...")
    t0 = time.time()
    verts, faces, norms, colors = tsdf_vol.get_mesh()
    t1 = time.time()
    fusion.meshwrite(os.path.join(args.out, "mesh.ply"), verts, faces, norms, colors)
    print("  %d vertices, %d faces; get_mesh %.1f ms, meshwrite %.1f ms" % (len(verts), len(faces), 1e3 * (t1 - t0),
                                                                          1e3 * (time.time() - t1)))

    print("Saving point cloud to pc.ply...")
    point_cloud = tsdf_vol.get_point_cloud()
    fusion.pcwrite(os.path.join(args.out, "pc.ply"), point_cloud)
