"""GPU parity of TSDFVolume.integrate against the oracle (restatement of fusion.py:260-303).

Gates (BASELINE.json north_star / SURVEY.md §8(d)): tsdf and weight max|d| <= 1e-5, colour plane bit-exact.
"""
import numpy as np
import pytest

from hive_b200.synthetic import Scene

pytestmark = pytest.mark.gpu

TOL = 1e-5


def _run_pair(scene, n_frames, c_oracle, obs_weight=1.0, frames=None, kernel=0):
    from hive_b200.fusion import TSDFVolume
    gpu = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    gpu.count_updates = True
    gpu.integrate_kernel = kernel
    orc = c_oracle.COracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    total = 0
    for i in range(n_frames):
        fr = scene.frame(i) if frames is None else frames[i]
        gpu.integrate(fr.rgb, fr.depth, scene.K, fr.pose, obs_weight)
        orc.integrate(fr.rgb, fr.depth, scene.K, fr.pose, obs_weight)
        total += orc.last_updated
    return gpu, orc, total


def _check(gpu, orc, total):
    tsdf, color = gpu.get_volume()
    weight = gpu.get_weight_volume()
    assert tsdf.shape == tuple(orc._vol_dim) and tsdf.dtype == np.float32
    assert color.shape == tuple(orc._vol_dim) and color.dtype == np.float32
    assert gpu.num_updated() == total
    assert np.array_equal(weight, orc._weight_vol_cpu)
    assert np.abs(tsdf - orc._tsdf_vol_cpu).max() <= TOL
    assert np.array_equal(color, orc._color_vol_cpu)  # bit-exact after the reference's packing
    return float(np.abs(tsdf - orc._tsdf_vol_cpu).max())


@pytest.mark.parametrize("kernel", [0, 1, 2, 3, 4])  # include/hive_b200.h HV_INTEGRATE_KERNEL_*
@pytest.mark.parametrize("variant", ["room", "dense"])
@pytest.mark.parametrize("config", ["tiny", "small"])
def test_integrate_matches_oracle(config, variant, kernel, c_oracle):
    scene = Scene(config, variant)
    gpu, orc, total = _run_pair(scene, scene.num_frames, c_oracle, kernel=kernel)
    assert total > 0
    _check(gpu, orc, total)


def test_integrate_c1_full_size(c_oracle):
    """BASELINE config C1 geometry (256^3, 640x480), a few frames."""
    scene = Scene("C1", "room")
    gpu, orc, total = _run_pair(scene, 3, c_oracle)
    _check(gpu, orc, total)


def test_integrate_many_frames_accumulation(c_oracle):
    """Running mean over many observations of the same voxels stays within tolerance (no drift)."""
    scene = Scene("tiny", "dense")
    frames = [scene.frame(i % 4) for i in range(120)]
    gpu, orc, total = _run_pair(scene, 120, c_oracle, frames=frames)
    _check(gpu, orc, total)


@pytest.mark.parametrize("kernel", [1, 2, 3])
@pytest.mark.parametrize("obs_weight", [2.0, 0.3])
def test_integrate_obs_weight(obs_weight, kernel, c_oracle):
    scene = Scene("tiny", "room")
    gpu, orc, total = _run_pair(scene, 4, c_oracle, obs_weight=obs_weight, kernel=kernel)
    _check(gpu, orc, total)


def test_integrate_identity_pose_ties(c_oracle):
    """Axis-aligned camera: pixel coordinates hit exact .5 ties (cx = 319.5), which must round half-to-even."""
    from hive_b200.synthetic import Frame
    scene = Scene("tiny", "room")
    fr = scene.frame(0)
    pose = np.eye(4)
    pose[:3, 3] = [0.0, 0.0, 0.5]
    frames = [Frame(fr.rgb, fr.depth, fr.mask, pose)] * 2
    gpu, orc, total = _run_pair(scene, 2, c_oracle, frames=frames)
    _check(gpu, orc, total)


def test_integrate_empty_and_invalid_frames(c_oracle):
    """All-zero depth updates nothing; a camera looking away updates nothing; the volume stays at its initial state."""
    from hive_b200.fusion import TSDFVolume
    scene = Scene("tiny", "room")
    fr = scene.frame(0)
    vol = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    vol.count_updates = True
    vol.integrate(fr.rgb, np.zeros_like(fr.depth), scene.K, fr.pose)
    away = fr.pose.copy()
    away[:3, :3] = away[:3, :3] @ np.diag([1.0, -1.0, -1.0])
    away[:3, 3] = [0, 0, -50.0]
    vol.integrate(fr.rgb, fr.depth, scene.K, away)
    tsdf, color = vol.get_volume()
    assert vol.num_updated() == 0
    assert (tsdf == 1.0).all() and (color == 0.0).all() and (vol.get_weight_volume() == 0.0).all()


def test_integrate_float64_inputs_like_demo(c_oracle):
    """demo.py:46-50 feeds float64 depth / intrinsics and non-contiguous colour; both paths see float32 depth."""
    scene = Scene("tiny", "room")
    from hive_b200.fusion import TSDFVolume
    gpu = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    orc = c_oracle.COracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for i in range(3):
        fr = scene.frame(i)
        bgr = fr.rgb[..., ::-1].copy()
        gpu.integrate(bgr[..., ::-1], fr.depth.astype(np.float64), scene.K.astype(np.float64), fr.pose)
        orc.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
    tsdf, color = gpu.get_volume()
    assert np.abs(tsdf - orc._tsdf_vol_cpu).max() <= TOL
    assert np.array_equal(color, orc._color_vol_cpu)


@pytest.mark.parametrize("kernel", [3])
def test_integrate_culling_grazing_poses(kernel, c_oracle):
    """Segment-level frustum culling must never drop a voxel the oracle updates: cameras whose image borders graze
    the volume at shallow angles, cameras inside the volume (segments crossing the camera plane), large rotations."""
    from hive_b200.synthetic import Frame
    scene = Scene("small", "room")
    rng = np.random.default_rng(7)
    base = [scene.frame(i) for i in range(4)]
    frames = []
    for i in range(12):
        fr = base[i % 4]
        a = rng.uniform(-np.pi, np.pi, 3) * (0.15 if i < 4 else 1.0)
        cx, sx, cy, sy, cz, sz = np.cos(a[0]), np.sin(a[0]), np.cos(a[1]), np.sin(a[1]), np.cos(a[2]), np.sin(a[2])
        R = (np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]) @ np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]]) @
             np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]))
        pose = fr.pose.copy()
        pose[:3, :3] = pose[:3, :3] @ R
        pose[:3, 3] += rng.uniform(-1.5, 1.5, 3)
        frames.append(Frame(fr.rgb, fr.depth, fr.mask, pose))
    gpu, orc, total = _run_pair(scene, len(frames), c_oracle, frames=frames, kernel=kernel)
    assert total > 0
    _check(gpu, orc, total)


@pytest.mark.parametrize("kernel", [1, 2, 3])
def test_integrate_ragged_z(kernel, c_oracle):
    """Z not a multiple of 4 or 128: the last float4 of a column is partial and the last segment is short."""
    from hive_b200 import synthetic
    synthetic.CONFIGS["ragged"] = dict(dim=(33, 29, 131), vs=0.04, frames=4, H=60, W=80)
    scene = Scene("ragged", "room")
    gpu, orc, total = _run_pair(scene, 4, c_oracle, kernel=kernel)
    assert total > 0
    _check(gpu, orc, total)


def test_integrate_matches_reference_golden_on_7scenes_data():
    """GPU against planes produced by the reference itself on its own 7-Scenes fixture frames (float64 depth and
    intrinsics, real poses, real sensor noise): tests/golden/tsdf_7scenes.npz, see oracle/make_golden.py."""
    import os
    from hive_b200.fusion import TSDFVolume
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tsdf_7scenes.npz"))
    vol = TSDFVolume(g["vol_bnds"].copy(), float(g["voxel_size"]))
    for d, c, p in zip(g["depth"], g["rgb"], g["pose"]):
        vol.integrate(c, d, g["cam_intr"], p, obs_weight=1.)
    tsdf, color = vol.get_volume()
    assert tsdf.shape == tuple(g["vol_dim"]) and np.array_equal(vol._vol_origin, g["vol_origin"])
    assert np.array_equal(vol.get_weight_volume(), g["weight"])
    assert np.abs(tsdf - g["tsdf"]).max() <= TOL
    assert np.array_equal(color, g["color"])
