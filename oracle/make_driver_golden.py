"""Golden vectors of the reference's own TSDF driver and key-frame selection (build container only).

    python -m oracle.make_driver_golden

TEST INFRASTRUCTURE.  Runs the *reference's* ``adjust_voxel_size`` / ``tsdf_fusion`` (hive/fusion.py:37-134, over its own
CPU ``TSDFVolume`` and its own cv2 ``dilate_mask``) and ``HiveDataset.select_key_frames`` (hive/io.py:1117-1189), compiled
from the source files by ``oracle/ref_shim.py``, on the seeded synthetic dataset.  Output: ``tests/golden/driver_reference.npz``

``fu<i>_*``  driver runs on ``small``/``room``, frames 0, 2, 3, 5, three dilation iterations, 6 cm voxels: (0) masks applied,
             no voxel budget; (1) masks applied, budget 300 000 voxels (voxel size raised); (2) inpainted data (no masks).
             Stored: voxel size, bounds as returned, and the three planes of the reference volume after the loop.
``mp_*``     the reference's ``get_mesh`` / ``get_point_cloud`` post-processing (fusion.py:311-350) of run 0's volume, with the
             absent scikit-image call answered by the C restatement of marching cubes
``kf_*``     key-frame lists of a 12-frame panning trajectory for thresholds 0.2-0.9 and frame steps 1 and 3; the
             world-to-camera matrices are stored as the reference derives them from its 7-vectors (pose_vec2mat).
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from hive_b200.synthetic import Scene, SyntheticDataset, panning_poses  # noqa: E402
from oracle import ref_shim  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
FRAME_SET = [0, 2, 3, 5]
FUSION_CASES = [(None, False), (300_000, False), (None, True)]  # (sdf_max_voxels, has_inpainted_frame_data)
KF_THRESHOLDS = [0.2, 0.3, 0.5, 0.7, 0.9]
KF_STEPS = [1, 3]
KF_FRAMES = 12


def fusion_cases(out):
    fusion = ref_shim.reference_fusion()
    created = []

    class Spy(fusion.TSDFVolume):
        def __init__(self, vol_bnds, voxel_size, use_gpu=False):
            super().__init__(vol_bnds, voxel_size, use_gpu=False)
            created.append(self)

        def get_mesh(self):  # needs scikit-image, absent; the planes are what is pinned
            return None, None, None, None

    ref = ref_shim.reference_tsdf_driver(Spy)
    sc = Scene("small", "room")
    for i, (max_voxels, inpainted) in enumerate(FUSION_CASES):
        options = ref["BackgroundMeshOptions"](depth_mask_dilation_iterations=3, sdf_voxel_size=0.06,
                                               sdf_max_voxels=max_voxels)
        voxel_size, bounds = ref["adjust_voxel_size"](SyntheticDataset(sc, 6, inpainted), options, FRAME_SET)
        del created[:]
        ref["tsdf_fusion"](SyntheticDataset(sc, 6, inpainted), options, frame_set=FRAME_SET)
        vol = created[0]
        out[f"fu{i}_voxel_size"] = np.float64(voxel_size)
        out[f"fu{i}_bounds"] = bounds
        out[f"fu{i}_max_voxels"] = np.int64(max_voxels or 0)
        out[f"fu{i}_inpainted"] = np.bool_(inpainted)
        out[f"fu{i}_tsdf"] = vol._tsdf_vol_cpu
        out[f"fu{i}_weight"] = vol._weight_vol_cpu
        out[f"fu{i}_color"] = vol._color_vol_cpu
        print(f"fusion case {i}: voxel {voxel_size:.4f}, dims {vol._vol_dim}, "
              f"{int((vol._weight_vol_cpu > 0).sum())} voxels observed")
        if i == 0:
            mesh_post_case(out, vol, fusion)
    out["fu_cases"] = len(FUSION_CASES)
    out["fu_frame_set"] = np.array(FRAME_SET)


def mesh_post_case(out, vol, fusion):
    """The reference's own ``get_mesh`` / ``get_point_cloud`` (fusion.py:311-350) on the volume of driver run 0.  The
    marching cubes inside them is scikit-image's (absent): ``measure.marching_cubes`` is answered by the C restatement, so
    what this pins is everything the reference does *after* it -- voxel rounding, world transform, colour gather and
    unpacking, the (V, 6) point-cloud layout -- in the reference's own arithmetic and dtypes."""
    from oracle import c_oracle
    c_oracle.build()
    fusion.measure.marching_cubes = lambda tsdf, level=0: c_oracle.marching_cubes(tsdf, float(level))
    fusion.measure.marching_cubes_lewiner = fusion.measure.marching_cubes
    vol.ctx = types.SimpleNamespace(pop=lambda: None)  # get_mesh pops the PyCUDA context even in CPU mode (fusion.py:349)
    verts, faces, norms, colors = fusion.TSDFVolume.get_mesh(vol)
    pc = fusion.TSDFVolume.get_point_cloud(vol)
    out.update(mp_verts=verts, mp_faces=faces, mp_norms=norms, mp_colors=colors, mp_pc=pc)
    print(f"get_mesh: verts {verts.dtype}{verts.shape}, faces {faces.dtype}{faces.shape}, colors {colors.dtype}, "
          f"point cloud {pc.dtype}{pc.shape}")


def key_frame_cases(out):
    geo = ref_shim.reference_geometric()
    select = ref_shim.reference_select_key_frames()
    sc = Scene("small", "room")
    ds = SyntheticDataset(sc, KF_FRAMES)
    vecs = [geo.pose_mat2vec(m) for m in panning_poses(sc, KF_FRAMES, 9.0)]  # world-to-camera 7-vectors, as HIVE stores them
    mats = np.stack([geo.pose_vec2mat(v) for v in vecs])
    h, w = sc.frame(0).depth.shape
    me = types.SimpleNamespace(num_frames=KF_FRAMES, metadata=types.SimpleNamespace(width=w, height=h),
                               camera_matrix=ds.camera_matrix, bg_depth_dataset=ds.bg_depth_dataset,
                               mask_dataset=ds.mask_dataset, camera_trajectory=vecs)
    out["kf_world_to_camera"] = mats
    out["kf_thresholds"] = np.array(KF_THRESHOLDS)
    out["kf_steps"] = np.array(KF_STEPS)
    for a, thr in enumerate(KF_THRESHOLDS):
        for b, step in enumerate(KF_STEPS):
            keys = select(me, threshold=thr, frame_step=step)
            out[f"kf_{a}_{b}"] = np.array(keys, dtype=np.int64)
            print(f"key frames thr {thr} step {step}: {keys}")


def main():
    assert ref_shim.available(), "the reference tree is needed to generate golden vectors"
    out = {}
    fusion_cases(out)
    key_frame_cases(out)
    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, "driver_reference.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
