"""Drop-in for the TSDF driver of ``hive/fusion.py`` (lines 37-134): ``adjust_voxel_size`` and ``tsdf_fusion``.

Same signatures, same host arithmetic for the volume bounds and the voxel budget, same frame loop -- but the volume
lives in B200 HBM (``hive_b200.fusion.TSDFVolume``), the per-frame ``dilate_mask`` + ``depth_im[mask > 0] = 0``
(hive/fusion.py:118-121) is fused into the frame upload instead of running cv2 on the host, and the mesh comes from the
CUDA marching cubes.  ``dataset`` is duck-typed on what the reference reads from a ``HiveDataset``:
``num_frames``, ``camera_matrix``, ``camera_trajectory.inverse().to_homogenous_transforms()``, ``bg_rgb_dataset[i]``,
``bg_depth_dataset[i]``, ``mask_dataset[i]``, ``has_inpainted_frame_data``.

The reference returns a ``trimesh.Trimesh``; trimesh is used when it is importable, otherwise a ``Mesh`` record with the
same four arrays under the same attribute names.
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

from . import fusion

__all__ = ["BackgroundMeshOptions", "Mesh", "adjust_voxel_size", "tsdf_fusion"]


@dataclass
class BackgroundMeshOptions:
    """The fields of hive/options.py:357-359 that the TSDF driver reads (same defaults)."""
    depth_mask_dilation_iterations: int = 10
    sdf_volume_size: float = 5.0
    sdf_voxel_size: float = 0.005
    sdf_max_voxels: Optional[int] = 320_000_000
    key_frame_threshold: float = 0.3
    key_frame_step: int = 30


@dataclass
class Mesh:
    """Stand-in for ``trimesh.Trimesh(vertices, faces, vertex_colors, vertex_normals)`` (hive/fusion.py:132)."""
    vertices: np.ndarray
    faces: np.ndarray
    vertex_colors: np.ndarray
    vertex_normals: np.ndarray


def _poses(dataset):
    # "poses are world-to-cam, but TSDFFusion appears to expect cam-to-world poses" (hive/fusion.py:50-51, 110-111)
    return dataset.camera_trajectory.inverse().to_homogenous_transforms()


def adjust_voxel_size(dataset, options: BackgroundMeshOptions, frame_set: List[int]) -> Tuple[float, np.ndarray]:
    """hive/fusion.py:37-76: AABB of all view frusta (always containing the origin) and the voxel-budget rule."""
    logging.info("Estimating voxel volume bounds...")
    vol_bnds = np.zeros((3, 2))
    camera_trajectory = _poses(dataset)
    for i in frame_set:
        depth_im = dataset.bg_depth_dataset[i]
        cam_pose = camera_trajectory[i]
        view_frust_pts = fusion.get_view_frustum(depth_im, dataset.camera_matrix, cam_pose)
        vol_bnds[:, 0] = np.minimum(vol_bnds[:, 0], np.amin(view_frust_pts, axis=1))
        vol_bnds[:, 1] = np.maximum(vol_bnds[:, 1], np.amax(view_frust_pts, axis=1))

    extent = vol_bnds[:, 1] - vol_bnds[:, 0]
    voxel_count = np.ceil(np.prod(extent / options.sdf_voxel_size))
    if options.sdf_max_voxels and voxel_count > options.sdf_max_voxels:
        voxel_size = (np.prod(extent) / options.sdf_max_voxels) ** (1 / 3)
        logging.info(f"Increasing voxel size to {voxel_size:.3f}: Using a voxel size of {options.sdf_voxel_size} would "
                     f"result in {voxel_count:,.0f} voxels, which is above the specified limit of "
                     f"{options.sdf_max_voxels:,d}.")
    else:
        voxel_size = options.sdf_voxel_size
    return voxel_size, vol_bnds


def tsdf_fusion(dataset, options: BackgroundMeshOptions = None, num_frames=-1, frame_set: Optional[List[int]] = None,
                srgb_vertex_colors=False):
    """hive/fusion.py:79-134: fuse the (masked) frames of ``dataset`` and return the marching-cubes mesh.

    ``srgb_vertex_colors=True`` folds in what the only caller does to the result next (hive/pipeline.py:281-282,
    ``vertex_colors[:, :3] = (255 * np.power(vertex_colors / 255, 2.2)).astype(int)``) while the colours are still in HBM.
    """
    options = options if options is not None else BackgroundMeshOptions()
    if num_frames == -1:
        num_frames = dataset.num_frames
    if frame_set is None:
        frame_set = range(num_frames)

    voxel_size, volume_bounds = adjust_voxel_size(dataset=dataset, options=options, frame_set=frame_set)
    logging.info("Initializing voxel volume...")
    tsdf_vol = fusion.TSDFVolume(volume_bounds, voxel_size=voxel_size)

    logging.info("Fusing frames...")
    has_inpainted_frame_data = dataset.has_inpainted_frame_data
    camera_trajectory = _poses(dataset)
    for i in frame_set:
        color_image = dataset.bg_rgb_dataset[i]
        depth_im = dataset.bg_depth_dataset[i]
        cam_pose = camera_trajectory[i]
        if not has_inpainted_frame_data:
            # dilate + zero on the device, inside the upload (the reference: cv2.dilate on the host, then fancy indexing)
            tsdf_vol.integrate(color_image, depth_im, dataset.camera_matrix, cam_pose, obs_weight=1.,
                               mask=dataset.mask_dataset[i],
                               mask_dilation_iterations=options.depth_mask_dilation_iterations)
        else:
            tsdf_vol.integrate(color_image, depth_im, dataset.camera_matrix, cam_pose, obs_weight=1.)

    verts, faces, norms, colors = tsdf_vol.get_mesh(srgb_colors=srgb_vertex_colors)
    try:
        import trimesh
        return trimesh.Trimesh(vertices=verts, faces=faces, vertex_colors=colors, vertex_normals=norms)
    except ImportError:
        return Mesh(vertices=verts, faces=faces, vertex_colors=colors, vertex_normals=norms)


def select_key_frames(dataset, threshold=0.3, frame_step=30) -> List[int]:
    """``HiveDataset.select_key_frames`` (hive/io.py:1117-1189): frames whose static points, projected into every key
    frame so far, cover less than ``threshold`` of the image (bounding-box area of the visible projections).

    The reference materialises the candidate's point cloud and one projected copy per key frame on the host; here one
    kernel per candidate frame back-projects, projects into all current key frames and reduces the boxes
    (``hv_keyframe_overlap``), so a candidate costs one H2D of its depth + mask and 20 bytes per key frame back.
    ``dataset``: ``num_frames``, ``camera_matrix``, ``camera_trajectory.to_homogenous_transforms()`` (world-to-camera),
    ``bg_depth_dataset[i]``, ``mask_dataset[i]``.
    """
    import ctypes

    from . import _lib, _xfer, geometric
    logging.info(f"Selecting key frames (threshold={threshold})...")
    if not (0.0 <= threshold <= 1.0):
        raise ValueError(f"Threshold must be a real number between zero and one (inclusive), but got {threshold}.")
    if threshold == 0.0:
        return [0]
    elif threshold == 1.0:
        return list(range(dataset.num_frames))
    if frame_step < 1:
        raise ValueError(f"Frame step must be a positive integer, but got {frame_step} instead.")

    torch = _lib.require_cuda()
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    K = np.asarray(dataset.camera_matrix)
    kinv = geometric._kinv64(K)
    K64 = np.asarray(K, dtype=np.float64)
    poses = np.asarray(dataset.camera_trajectory.to_homogenous_transforms(), dtype=np.float64)

    def cam21(M, pose):
        return np.concatenate([M.reshape(9), pose[:3, :3].reshape(9), pose[:3, 3].reshape(3)])

    key_frames = [0]
    key_cams = [cam21(K64, poses[0])]
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for frame in range(1, dataset.num_frames, frame_step):
            depth = np.asarray(dataset.bg_depth_dataset[frame])
            height, width = depth.shape
            d = _xfer.to_device(depth, dev, np.float32)
            m = _xfer.to_device(np.asarray(dataset.mask_dataset[frame]) != 0, dev, np.uint8)  # 0 = static pixel
            cam = _xfer.to_device(cam21(kinv, poses[frame]), dev, np.float64)
            kc = _xfer.to_device(np.stack(key_cams), dev, np.float64)
            n_keys = len(key_frames)
            box = torch.empty((n_keys, 4), dtype=torch.int32, device=dev)
            count = torch.empty((n_keys,), dtype=torch.int32, device=dev)
            _lib.check(lib.hv_keyframe_overlap(ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(m.data_ptr()), height, width,
                                               ctypes.c_void_p(cam.data_ptr()), ctypes.c_void_p(kc.data_ptr()), n_keys,
                                               ctypes.c_void_p(box.data_ptr()), ctypes.c_void_p(count.data_ptr()), stream))
            box_h, count_h = _xfer.to_host(box, count)
            for k in range(n_keys):  # io.py:1157-1185, in key-frame order
                if count_h[k] == 0:
                    continue
                lo, hi = box_h[k, :2].astype(np.int64), box_h[k, 2:].astype(np.int64)
                overlap_ratio = np.prod(hi - lo) / (width * height)
                if overlap_ratio >= threshold:
                    break
            else:
                key_frames.append(frame)
                key_cams.append(cam21(K64, poses[frame]))
    return key_frames
