"""CPU: pins the driver restatements (oracle/driver_oracle.py) against the reference's own ``adjust_voxel_size`` /
``tsdf_fusion`` (hive/fusion.py:37-134) and ``HiveDataset.select_key_frames`` (hive/io.py:1117-1189), run in the build
container by oracle/make_driver_golden.py -> tests/golden/driver_reference.npz."""
import os

import numpy as np
import pytest

from hive_b200.synthetic import Scene, SyntheticDataset, _Trajectory
from oracle import driver_oracle as drv

GOLD = os.path.join(os.path.dirname(__file__), "golden", "driver_reference.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.mark.parametrize("case", [0, 1, 2])
def test_tsdf_fusion_restatement_equals_the_reference_driver(gold, case, c_oracle):
    sc = Scene("small", "room")
    max_voxels = int(gold[f"fu{case}_max_voxels"]) or None
    inpainted = bool(gold[f"fu{case}_inpainted"])
    frame_set = [int(i) for i in gold["fu_frame_set"]]
    vs, bounds = drv.adjust_voxel_size(SyntheticDataset(sc, 6, inpainted), 0.06, max_voxels, frame_set)
    assert vs == float(gold[f"fu{case}_voxel_size"])
    assert np.array_equal(bounds, gold[f"fu{case}_bounds"])
    vol, _ = drv.tsdf_fusion(SyntheticDataset(sc, 6, inpainted), 0.06, max_voxels, 3, frame_set)
    assert vol._tsdf_vol_cpu.shape == gold[f"fu{case}_tsdf"].shape
    assert np.abs(vol._tsdf_vol_cpu - gold[f"fu{case}_tsdf"]).max() <= 1e-6
    assert np.array_equal(vol._weight_vol_cpu, gold[f"fu{case}_weight"])
    assert np.array_equal(vol._color_vol_cpu, gold[f"fu{case}_color"])
    assert (gold[f"fu{case}_weight"] > 0).sum() > 50_000


def key_frame_dataset(gold):
    mats = gold["kf_world_to_camera"]
    ds = SyntheticDataset(Scene("small", "room"), len(mats))
    ds.camera_trajectory = _Trajectory(list(mats))
    return ds


def test_select_key_frames_restatement_equals_the_reference(gold):
    ds = key_frame_dataset(gold)
    seen = set()
    for a, thr in enumerate(gold["kf_thresholds"]):
        for b, step in enumerate(gold["kf_steps"]):
            want = [int(k) for k in gold[f"kf_{a}_{b}"]]
            assert drv.select_key_frames(ds, float(thr), int(step)) == want
            seen.add(tuple(want))
    assert len(seen) >= 6  # the cases really differ


def test_mesh_postprocess_restatement_equals_the_reference_get_mesh(gold, c_oracle):
    """fusion.py:339-348 / :318-329 (everything after the marching-cubes call) in the reference's own arithmetic: the
    golden was produced by the reference's get_mesh / get_point_cloud with the C marching cubes standing in for skimage."""
    from oracle import fusion_oracle as fo
    tsdf, color = gold["fu0_tsdf"], gold["fu0_color"]
    voxel_size = float(gold["fu0_voxel_size"])
    origin = gold["fu0_bounds"][:, 0].copy(order='C').astype(np.float32)  # fusion.py:46
    ov, of, on, _ = c_oracle.marching_cubes(tsdf, 0.0)
    verts, colors = fo.mesh_postprocess(ov, color, voxel_size, origin)
    assert verts.dtype == gold["mp_verts"].dtype == np.float32
    assert np.array_equal(verts, gold["mp_verts"]) and np.array_equal(colors, gold["mp_colors"])
    assert np.array_equal(of, gold["mp_faces"]) and np.array_equal(on, gold["mp_norms"])
    pc = np.hstack([verts, colors])
    assert pc.dtype == gold["mp_pc"].dtype and np.array_equal(pc, gold["mp_pc"])
    assert len(verts) > 10_000 and len(np.unique(colors)) > 50
