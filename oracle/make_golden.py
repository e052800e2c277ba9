"""Generate the golden vectors under tests/golden/ by running the *reference itself* (build container only).

    python -m oracle.make_golden

TEST INFRASTRUCTURE.  Imports the real ``third_party/tsdf_fusion_python/fusion.py`` (CPU mode, numpy + numba) and the
real ``hive/geometric.py`` from /root/reference through ``oracle/ref_shim.py`` and dumps their outputs on seeded
synthetic inputs (``hive_b200.synthetic``).  The fixtures travel to the GPU box, where /root/reference does not exist:

``tsdf_tiny_room.npz``   planes after 6 frames of the ``tiny``/``room`` scene (fusion.py:217-303, use_gpu=False)
``tsdf_tiny_dense.npz``  same for the ``dense`` variant, obs_weight 1 then 2
``frustum.npz``          get_view_frustum of three frames (fusion.py:361-375)
``backproject.npz``      point_cloud_from_depth / image2world / world2image of one frame (hive/geometric.py:107-206)

``tsdf_7scenes.npz``     the reference's own fixture data (third_party/tsdf_fusion_python/data, read as demo.py:19-55
                         reads it): frames 0/250/500/750/999 subsampled 4x (160x120, intrinsics scaled), bounds from the
                         reference's get_view_frustum, 6 cm voxels; inputs and the reference's output planes

Planes are stored losslessly (npz, deflate); every file also carries the SHA-256 of the raw little-endian bytes.
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from hive_b200.synthetic import Scene  # noqa: E402
from oracle import ref_shim  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def tsdf_case(fusion, variant: str, weights):
    scene = Scene("tiny", variant)
    vol = fusion.TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, use_gpu=False)
    for i in range(scene.num_frames):
        fr = scene.frame(i)
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose, obs_weight=weights[i % len(weights)])
    t, w, c = vol._tsdf_vol_cpu, vol._weight_vol_cpu, vol._color_vol_cpu
    return dict(tsdf=t, weight=w, color=c, vol_dim=np.asarray(vol._vol_dim), vol_origin=vol._vol_origin,
                weights=np.asarray(weights, dtype=np.float64), num_frames=scene.num_frames,
                sha_tsdf=sha(t), sha_weight=sha(w), sha_color=sha(c))


def seven_scenes_case(fusion, indices=(0, 250, 500, 750, 999), step=4, voxel_size=0.06):
    """The 7-Scenes frames shipped with the reference, read like demo.py:19-55, subsampled so the fixture stays small."""
    import cv2
    data = os.path.join(ref_shim.REFERENCE_ROOT, "third_party", "tsdf_fusion_python", "data")
    cam_intr = np.loadtxt(os.path.join(data, "camera-intrinsics.txt"), delimiter=' ')
    cam_intr[:2] /= step  # pixel (u, v) of the subsampled image = (u, v) / step of the full image
    depths, colors, poses = [], [], []
    vol_bnds = np.zeros((3, 2))
    for i in indices:
        depth_im = cv2.imread(os.path.join(data, "frame-%06d.depth.png" % i), -1).astype(float)
        depth_im /= 1000.
        depth_im[depth_im == 65.535] = 0
        depth_im = np.ascontiguousarray(depth_im[::step, ::step])
        color = cv2.cvtColor(cv2.imread(os.path.join(data, "frame-%06d.color.jpg" % i)), cv2.COLOR_BGR2RGB)
        color = np.ascontiguousarray(color[::step, ::step])
        pose = np.loadtxt(os.path.join(data, "frame-%06d.pose.txt" % i))
        pts = fusion.get_view_frustum(depth_im, cam_intr, pose)
        vol_bnds[:, 0] = np.minimum(vol_bnds[:, 0], np.amin(pts, axis=1))
        vol_bnds[:, 1] = np.maximum(vol_bnds[:, 1], np.amax(pts, axis=1))
        depths.append(depth_im)
        colors.append(color)
        poses.append(pose)
    bounds_in = vol_bnds.copy()
    vol = fusion.TSDFVolume(vol_bnds, voxel_size, use_gpu=False)
    for d, c, p in zip(depths, colors, poses):
        vol.integrate(c, d, cam_intr, p, obs_weight=1.)
    t, w, c = vol._tsdf_vol_cpu, vol._weight_vol_cpu, vol._color_vol_cpu
    return dict(depth=np.stack(depths), rgb=np.stack(colors), pose=np.stack(poses), cam_intr=cam_intr,
                vol_bnds=bounds_in, voxel_size=voxel_size, indices=np.asarray(indices), tsdf=t, weight=w, color=c,
                vol_dim=np.asarray(vol._vol_dim), vol_origin=vol._vol_origin, sha_tsdf=sha(t), sha_weight=sha(w),
                sha_color=sha(c))


def main():
    assert ref_shim.available(), "the reference tree is needed to generate golden vectors"
    os.makedirs(GOLDEN, exist_ok=True)
    fusion = ref_shim.reference_fusion()
    np.savez_compressed(os.path.join(GOLDEN, "tsdf_tiny_room.npz"), **tsdf_case(fusion, "room", [1.0]))
    np.savez_compressed(os.path.join(GOLDEN, "tsdf_tiny_dense.npz"), **tsdf_case(fusion, "dense", [1.0, 2.0]))

    np.savez_compressed(os.path.join(GOLDEN, "tsdf_7scenes.npz"), **seven_scenes_case(fusion))

    scene = Scene("tiny", "room")
    fr = [scene.frame(i) for i in range(3)]
    np.savez_compressed(os.path.join(GOLDEN, "frustum.npz"),
                        **{f"f{i}": fusion.get_view_frustum(f.depth, scene.K, f.pose) for i, f in enumerate(fr)})

    geo = ref_shim.reference_geometric()
    scene = Scene("small", "room")
    f = scene.frame(2)
    w2c = np.linalg.inv(f.pose)  # hive poses are world-to-camera (hive/pipeline.py:357)
    R, t = w2c[:3, :3], w2c[:3, 3:]
    mask = f.mask > 0
    pts = geo.point_cloud_from_depth(f.depth, mask, scene.K, R, t)
    pts_nomask = geo.point_cloud_from_depth(f.depth, np.ones_like(mask), scene.K, R, t)
    uv, d = geo.world2image(pts, scene.K, R, t)
    np.savez_compressed(os.path.join(GOLDEN, "backproject.npz"), frame_index=2, R=R, t=t, points=pts,
                        points_nomask_sha=sha(pts_nomask), points_nomask_head=pts_nomask[:64],
                        n_nomask=len(pts_nomask), uv=uv, depth=d)
    for name in sorted(os.listdir(GOLDEN)):
        print(name, os.path.getsize(os.path.join(GOLDEN, name)))


if __name__ == "__main__":
    main()
