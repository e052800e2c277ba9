"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the TSDF driver around the volume (never imported by hive_b200/).

Restates, function by function:
  * ``dilate_mask``        hive/image_processing.py:32-48 (cv2.dilate, 3x3 rectangle of hive/options.py:248, n iterations,
                           then astype(bool)).  Two forms: ``dilate_mask_cv2`` calls OpenCV exactly as the reference does
                           (cv2 4.13 is in this image; the reference pins no cv2 version), ``dilate_mask_numpy`` is the
                           closed form "any non-zero pixel within Chebyshev distance n" used when cv2 is absent;
                           tests/test_oracle_driver.py pins the closed form against cv2.
  * ``adjust_voxel_size``  hive/fusion.py:37-76
  * ``tsdf_fusion``        hive/fusion.py:79-134, over the C oracle volume (oracle/tsdf_oracle.c) and the C marching cubes.
"""
from __future__ import annotations

import numpy as np

from . import fusion_oracle as fo


def dilate_mask_numpy(mask, num_iterations: int):
    m = np.asarray(mask).astype(np.float32) != 0
    n = int(num_iterations)
    if n == 0:
        return m
    H, W = m.shape
    pad = np.zeros((H + 2 * n, W + 2 * n), dtype=bool)
    pad[n:n + H, n:n + W] = m
    rows = np.zeros((H + 2 * n, W), dtype=bool)
    for s in range(2 * n + 1):  # horizontal pass
        rows |= pad[:, s:s + W]
    out = np.zeros((H, W), dtype=bool)
    for s in range(2 * n + 1):  # vertical pass
        out |= rows[s:s + H]
    return out


def dilate_mask_cv2(mask, num_iterations: int):
    import cv2
    kernel = cv2.getStructuringElement(cv2.MORPH_RECT, (3, 3))  # hive/options.py:248
    m = np.asarray(mask).astype(np.float32)
    m = cv2.dilate(m.astype(float), kernel, iterations=int(num_iterations))  # image_processing.py:41-42
    return m.astype(bool)


def dilate_mask(mask, num_iterations: int):
    try:
        return dilate_mask_cv2(mask, num_iterations)
    except ImportError:
        return dilate_mask_numpy(mask, num_iterations)


def adjust_voxel_size(dataset, sdf_voxel_size, sdf_max_voxels, frame_set):
    vol_bnds = np.zeros((3, 2))
    poses = dataset.camera_trajectory.inverse().to_homogenous_transforms()
    for i in frame_set:
        pts = fo.get_view_frustum(dataset.bg_depth_dataset[i], dataset.camera_matrix, poses[i])
        vol_bnds[:, 0] = np.minimum(vol_bnds[:, 0], np.amin(pts, axis=1))
        vol_bnds[:, 1] = np.maximum(vol_bnds[:, 1], np.amax(pts, axis=1))
    voxel_count = np.ceil(np.prod((vol_bnds[:, 1] - vol_bnds[:, 0]) / sdf_voxel_size))
    if sdf_max_voxels and voxel_count > sdf_max_voxels:
        voxel_size = (np.prod(vol_bnds[:, 1] - vol_bnds[:, 0]) / sdf_max_voxels) ** (1 / 3)
    else:
        voxel_size = sdf_voxel_size
    return voxel_size, vol_bnds


def tsdf_fusion(dataset, sdf_voxel_size=0.005, sdf_max_voxels=320_000_000, dilation_iterations=10, frame_set=None):
    """Returns (volume oracle object, (verts, faces, norms, colors))."""
    from . import c_oracle
    if frame_set is None:
        frame_set = range(dataset.num_frames)
    voxel_size, bounds = adjust_voxel_size(dataset, sdf_voxel_size, sdf_max_voxels, frame_set)
    vol = c_oracle.COracleTSDFVolume(bounds, voxel_size)
    poses = dataset.camera_trajectory.inverse().to_homogenous_transforms()
    for i in frame_set:
        color = dataset.bg_rgb_dataset[i]
        depth = dataset.bg_depth_dataset[i]
        if not dataset.has_inpainted_frame_data:
            mask = dilate_mask(dataset.mask_dataset[i], dilation_iterations)
            depth[mask > 0] = 0.0  # hive/fusion.py:121 (in place on the array the dataset handed out)
        vol.integrate(color, depth, dataset.camera_matrix, poses[i], obs_weight=1.0)
    return vol, c_oracle.get_mesh(vol._tsdf_vol_cpu, vol._color_vol_cpu, vol._voxel_size, vol._vol_origin)


def select_key_frames(dataset, threshold=0.3, frame_step=30):
    """hive/io.py:1117-1189 over the verbatim numpy restatements of hive.geometric (oracle/d2m_oracle.py, pinned against
    the real module by tests/golden/backproject.npz)."""
    from . import d2m_oracle as geo
    if not (0.0 <= threshold <= 1.0):
        raise ValueError(f"Threshold must be a real number between zero and one (inclusive), but got {threshold}.")
    if threshold == 0.0:
        return [0]
    elif threshold == 1.0:
        return list(range(dataset.num_frames))
    if frame_step < 1:
        raise ValueError(f"Frame step must be a positive integer, but got {frame_step} instead.")
    poses = dataset.camera_trajectory.to_homogenous_transforms()
    K = dataset.camera_matrix
    key_frames = [0]
    for frame in range(1, dataset.num_frames, frame_step):
        depth = dataset.bg_depth_dataset[frame]
        height, width = depth.shape
        mask = dataset.mask_dataset[frame] == 0
        R, t = poses[frame][:3, :3], poses[frame][:3, 3:]
        pts = geo.point_cloud_from_depth(depth, mask, K, R, t)
        for key_frame in key_frames:
            R, t = poses[key_frame][:3, :3], poses[key_frame][:3, 3:]
            proj, _ = geo.world2image(pts, K, R, t)
            vis = (proj[:, 0] >= 0) & (proj[:, 0] < width) & (proj[:, 1] >= 0) & (proj[:, 1] < height)
            visible = proj[vis]
            if len(visible) == 0:
                continue
            area = np.prod(np.max(visible, axis=0) - np.min(visible, axis=0))
            if area / (width * height) >= threshold:
                break
        else:
            key_frames.append(frame)
    return key_frames


def vertex_colors_to_srgb(vertex_colors):
    """hive/pipeline.py:281-282, on a copy: ``vertex_colors[:, :3] = (255 * np.power(vertex_colors / 255, 2.2)).astype(int)``
    (float64 power, truncation, stored back into the uint8 colour array)."""
    out = np.array(vertex_colors, copy=True)
    rgb = out[:, :3]
    out[:, :3] = (255 * np.power(rgb / 255, 2.2)).astype(int)
    return out
