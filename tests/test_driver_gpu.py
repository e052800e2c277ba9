"""GPU parity of the TSDF driver pieces around the volume: dilate_mask
(hive/image_processing.py:32-48), the masked frame upload (hive/fusion.py:118-121), per-frame depth maxima
(fusion.py:364) and the whole ``tsdf_fusion`` loop (hive/fusion.py:79-134) against oracle/driver_oracle.py.
Bit-exact for masks; tsdf <= 1e-5, weight and colour bit-exact; mesh identical to the oracle's marching cubes."""
import numpy as np
import pytest

from hive_b200.synthetic import Scene, SyntheticDataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(60, 80), (37, 45), (33, 64), (5, 3), (480, 640)])
@pytest.mark.parametrize("n", [0, 1, 2, 10, 40])
def test_dilate_mask_matches_oracle(shape, n):
    from hive_b200.image_processing import MaskDilationOptions, dilate_mask
    from oracle import driver_oracle as drv
    rng = np.random.default_rng(11 + n)
    mask = (rng.random(shape) < 0.01).astype(np.uint8) * rng.integers(1, 200, size=shape).astype(np.uint8)
    got = dilate_mask(mask, MaskDilationOptions(num_iterations=n))
    assert got.dtype == np.bool_ and got.shape == shape
    assert np.array_equal(got, drv.dilate_mask(mask, n))
    # bool and float masks go through the same "non-zero" rule (image_processing.py:41-43)
    assert np.array_equal(dilate_mask(mask > 0, MaskDilationOptions(n)), got)
    assert np.array_equal(dilate_mask(mask.astype(np.float64) * 0.25, MaskDilationOptions(n)), got)


def test_dilate_mask_edge_cases():
    from hive_b200.image_processing import MaskDilationOptions, dilate_mask
    from oracle import driver_oracle as drv
    opts = MaskDilationOptions(num_iterations=3)
    assert not dilate_mask(np.zeros((20, 31), np.uint8), opts).any()
    assert dilate_mask(np.ones((20, 31), np.uint8), opts).all()
    corner = np.zeros((20, 70), np.uint8)
    corner[0, 0] = corner[19, 69] = 2
    assert dilate_mask(corner, opts).sum() == 2 * 16
    with pytest.raises(AssertionError):
        dilate_mask(np.zeros((2, 3, 4), np.uint8), opts)
    # all-ones squares of odd size: n applications of k x k = one dilation by a square of radius n * (k // 2)
    rng = np.random.default_rng(5)
    sparse = (rng.random((37, 83)) < 0.01).astype(np.uint8)
    for k, n in ((5, 1), (5, 2), (7, 1), (1, 3)):
        got = dilate_mask(sparse, MaskDilationOptions(n, np.ones((k, k), np.uint8)))
        assert np.array_equal(got, drv.dilate_mask_numpy(sparse, n * (k // 2)))
    cross = np.array([[0, 1, 0], [1, 1, 1], [0, 1, 0]], np.uint8)  # cv2.MORPH_CROSS: not a rectangle
    for bad in (cross, np.ones((4, 4), np.uint8), np.ones((3, 5), np.uint8)):
        with pytest.raises(NotImplementedError):
            dilate_mask(corner, MaskDilationOptions(1, bad))


def test_dilate_masks_batched():
    import torch
    from hive_b200.image_processing import dilate_masks_device
    from oracle import driver_oracle as drv
    sc = Scene("small", "room")
    masks = np.stack([sc.frame(i).mask for i in range(4)])
    out = dilate_masks_device(torch.from_numpy(masks).cuda(), 10).cpu().numpy()
    for b in range(4):
        assert np.array_equal(out[b].astype(bool), drv.dilate_mask(masks[b], 10))


def test_depth_max():
    from hive_b200.image_processing import depth_max
    sc = Scene("small", "room")
    depths = np.stack([sc.frame(i).depth for i in range(5)])
    depths[3] = 0.0
    assert np.array_equal(depth_max(depths), depths.reshape(5, -1).max(axis=1))


def test_masked_integrate_equals_host_masking(c_oracle):
    """integrate(..., mask=, mask_dilation_iterations=) == the reference's host-side masking followed by integrate."""
    from hive_b200.fusion import TSDFVolume
    from oracle import driver_oracle as drv
    sc = Scene("small", "room")
    gpu = TSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
    orc = c_oracle.COracleTSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
    for i in range(4):
        fr = sc.frame(i)
        depth_before = fr.depth.copy()
        gpu.integrate(fr.rgb, fr.depth, sc.K, fr.pose, mask=fr.mask, mask_dilation_iterations=4)
        assert np.array_equal(fr.depth, depth_before)  # the caller's array is not modified
        d = fr.depth.copy()
        d[drv.dilate_mask(fr.mask, 4) > 0] = 0.0
        orc.integrate(fr.rgb, d, sc.K, fr.pose)
    tsdf, color = gpu.get_volume()
    assert np.abs(tsdf - orc._tsdf_vol_cpu).max() <= 1e-5
    assert np.array_equal(gpu.get_weight_volume(), orc._weight_vol_cpu)
    assert np.array_equal(color, orc._color_vol_cpu)


@pytest.mark.parametrize("max_voxels", [None, 300_000])
@pytest.mark.parametrize("inpainted", [False, True])
def test_tsdf_fusion_driver_matches_oracle(max_voxels, inpainted, c_oracle, monkeypatch):
    from hive_b200 import hive_fusion
    from oracle import driver_oracle as drv
    from oracle import fusion_oracle as fo
    sc = Scene("small", "room")
    frame_set = [0, 2, 3, 5]
    opts = hive_fusion.BackgroundMeshOptions(depth_mask_dilation_iterations=3, sdf_voxel_size=0.06,
                                             sdf_max_voxels=max_voxels)
    created = []
    base = hive_fusion.fusion.TSDFVolume

    class Spy(base):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            created.append(self)

    monkeypatch.setattr(hive_fusion.fusion, "TSDFVolume", Spy)
    mesh = hive_fusion.tsdf_fusion(SyntheticDataset(sc, 6, inpainted), opts, frame_set=frame_set)
    vol, _ = drv.tsdf_fusion(SyntheticDataset(sc, 6, inpainted), 0.06, max_voxels, 3, frame_set)
    vs, bounds = hive_fusion.adjust_voxel_size(SyntheticDataset(sc, 6), opts, frame_set)
    ovs, obounds = drv.adjust_voxel_size(SyntheticDataset(sc, 6), 0.06, max_voxels, frame_set)
    assert vs == ovs and np.array_equal(bounds, obounds)
    assert (vs > 0.06) == bool(max_voxels)

    # planes: the driver's volume against the oracle's driver
    gpu = created[0]
    tsdf, color = gpu.get_volume()
    assert tsdf.shape == vol._tsdf_vol_cpu.shape
    assert np.abs(tsdf - vol._tsdf_vol_cpu).max() <= 1e-5
    assert np.array_equal(gpu.get_weight_volume(), vol._weight_vol_cpu)
    assert np.array_equal(color, vol._color_vol_cpu)
    # mesh: the oracle's marching cubes on the very planes the GPU meshed (signs are then equal by construction)
    ov, of, on, _, ok = c_oracle.marching_cubes(tsdf, 0.0, with_keys=True)
    order = c_oracle.product_order(ok)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    ow, oc = fo.mesh_postprocess(ov[order], color, gpu._voxel_size, gpu._vol_origin)
    assert len(ov) > 1000
    assert np.array_equal(mesh.faces, rank[of].astype(np.int32))
    assert np.array_equal(mesh.vertices, ow) and np.array_equal(mesh.vertex_colors, oc)
    assert np.abs(mesh.vertex_normals - on[order]).max() <= 1e-6


def _rotating_dataset(n=12, step_deg=9.0):
    """Frames of the small room scene behind a camera that also pans `step_deg` per frame."""
    from hive_b200.synthetic import _Trajectory, panning_poses
    sc = Scene("small", "room")
    ds = SyntheticDataset(sc, n)
    ds.camera_trajectory = _Trajectory(panning_poses(sc, n, step_deg))
    return ds


@pytest.mark.parametrize("frame_step", [1, 3])
@pytest.mark.parametrize("threshold", [0.2, 0.3, 0.5, 0.7, 0.9])
def test_select_key_frames_matches_oracle(threshold, frame_step):
    """HiveDataset.select_key_frames (hive/io.py:1117-1189): same key-frame list as the numpy restatement."""
    from hive_b200 import hive_fusion
    from oracle import driver_oracle as drv
    ds = _rotating_dataset()
    want = drv.select_key_frames(ds, threshold, frame_step)
    assert hive_fusion.select_key_frames(ds, threshold, frame_step) == want
    assert 1 < len(want) <= 12


def test_select_key_frames_edge_cases():
    from hive_b200 import hive_fusion
    ds = _rotating_dataset(5)
    assert hive_fusion.select_key_frames(ds, 0.0) == [0]
    assert hive_fusion.select_key_frames(ds, 1.0) == [0, 1, 2, 3, 4]
    with pytest.raises(ValueError):
        hive_fusion.select_key_frames(ds, 1.5)
    with pytest.raises(ValueError):
        hive_fusion.select_key_frames(ds, 0.5, frame_step=0)
    # a fully masked candidate has no static point: it can never overlap, so it becomes a key frame (io.py:1169-1170)
    ds.mask_dataset = type(ds.mask_dataset)(lambda i: np.ones((ds.scene.H, ds.scene.W), np.uint8))
    assert hive_fusion.select_key_frames(ds, 0.5, frame_step=2) == [0, 1, 3]


def test_keyframe_boxes_match_oracle_projection():
    """The per-key-frame boxes and counts themselves, not only the decisions."""
    import ctypes
    import torch
    from hive_b200 import _lib, geometric
    from oracle import d2m_oracle as geo
    ds = _rotating_dataset(6)
    poses = ds.camera_trajectory.to_homogenous_transforms()
    K = ds.camera_matrix
    depth, mask = ds.bg_depth_dataset[4], ds.mask_dataset[4]
    H, W = depth.shape
    keys = [0, 2, 3]

    def cam21(M, p):
        return np.concatenate([np.asarray(M, np.float64).reshape(9), p[:3, :3].reshape(9), p[:3, 3].reshape(3)])
    dev = torch.device("cuda")
    d = torch.from_numpy(depth).to(dev)
    m = torch.from_numpy((mask != 0).astype(np.uint8)).to(dev)
    cam = torch.from_numpy(cam21(geometric._kinv64(K), poses[4])).to(dev)
    kc = torch.from_numpy(np.stack([cam21(K, poses[k]) for k in keys])).to(dev)
    box = torch.empty((3, 4), dtype=torch.int32, device=dev)
    cnt = torch.empty(3, dtype=torch.int32, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    _lib.check(_lib.load().hv_keyframe_overlap(p(d), p(m), H, W, p(cam), p(kc), 3, p(box), p(cnt),
                                               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    box, cnt = box.cpu().numpy(), cnt.cpu().numpy()
    pts = geo.point_cloud_from_depth(depth, mask == 0, K, poses[4][:3, :3], poses[4][:3, 3:])
    for j, k in enumerate(keys):
        uv, _ = geo.world2image(pts, K, poses[k][:3, :3], poses[k][:3, 3:])
        vis = (uv[:, 0] >= 0) & (uv[:, 0] < W) & (uv[:, 1] >= 0) & (uv[:, 1] < H)
        assert cnt[j] == vis.sum() > 0
        assert np.array_equal(box[j], np.concatenate([uv[vis].min(axis=0), uv[vis].max(axis=0)]))


@pytest.mark.parametrize("n_verts", [1, 5, 16, 1000, 100_003])
@pytest.mark.parametrize("cols", [3, 4])
def test_vertex_colors_to_srgb_matches_oracle(n_verts, cols):
    """hive/pipeline.py:281-282 through hv_byte_lut, ragged sizes around the 16-byte vector width; bit-exact."""
    from hive_b200 import image_processing
    from oracle import driver_oracle as drv
    rng = np.random.default_rng(n_verts + cols)
    c = rng.integers(0, 256, (n_verts, cols), dtype=np.uint8)
    before = c.copy()
    got = image_processing.vertex_colors_to_srgb(c)
    assert np.array_equal(c, before), "input must not be modified"
    assert got.dtype == np.uint8 and np.array_equal(got, drv.vertex_colors_to_srgb(c))


def test_get_mesh_srgb_colors_equals_conversion_of_plain_mesh():
    from hive_b200.fusion import TSDFVolume
    from oracle import driver_oracle as drv
    sc = Scene("small", "room")
    vol = TSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
    for i in range(3):
        fr = sc.frame(i)
        vol.integrate(fr.rgb, fr.depth, sc.K, fr.pose)
    v0, f0, n0, c0 = vol.get_mesh()
    v1, f1, n1, c1 = vol.get_mesh(srgb_colors=True)
    assert np.array_equal(v0, v1) and np.array_equal(f0, f1) and np.array_equal(n0, n1)
    assert len(c0) > 1000 and len(np.unique(c0)) > 50
    assert np.array_equal(c1, drv.vertex_colors_to_srgb(c0))
    # and through the driver
    from hive_b200 import hive_fusion
    opts = hive_fusion.BackgroundMeshOptions(depth_mask_dilation_iterations=2, sdf_voxel_size=0.06, sdf_max_voxels=None)
    plain = hive_fusion.tsdf_fusion(SyntheticDataset(sc, 4), opts)
    srgb = hive_fusion.tsdf_fusion(SyntheticDataset(sc, 4), opts, srgb_vertex_colors=True)
    assert np.array_equal(srgb.vertex_colors, drv.vertex_colors_to_srgb(plain.vertex_colors))
    assert np.array_equal(srgb.faces, plain.faces)


# ---- against the reference's own driver (tests/golden/driver_reference.npz, oracle/make_driver_golden.py) -------------

def _driver_gold():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "driver_reference.npz"))


@pytest.mark.parametrize("case", [0, 1, 2])
def test_tsdf_fusion_driver_matches_reference_golden(case, monkeypatch):
    """hive_b200.hive_fusion.tsdf_fusion (device dilation fused into the upload) against the planes the reference's own
    driver produced with cv2.dilate + its CPU TSDFVolume: voxel size and bounds equal, tsdf <= 1e-5, weight and colour
    bit-exact."""
    from hive_b200 import hive_fusion
    g = _driver_gold()
    sc = Scene("small", "room")
    max_voxels = int(g[f"fu{case}_max_voxels"]) or None
    inpainted = bool(g[f"fu{case}_inpainted"])
    frame_set = [int(i) for i in g["fu_frame_set"]]
    opts = hive_fusion.BackgroundMeshOptions(depth_mask_dilation_iterations=3, sdf_voxel_size=0.06,
                                             sdf_max_voxels=max_voxels)
    vs, bounds = hive_fusion.adjust_voxel_size(SyntheticDataset(sc, 6, inpainted), opts, frame_set)
    assert vs == float(g[f"fu{case}_voxel_size"]) and np.array_equal(bounds, g[f"fu{case}_bounds"])
    created = []
    base = hive_fusion.fusion.TSDFVolume

    class Spy(base):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            created.append(self)

    monkeypatch.setattr(hive_fusion.fusion, "TSDFVolume", Spy)
    hive_fusion.tsdf_fusion(SyntheticDataset(sc, 6, inpainted), opts, frame_set=frame_set)
    tsdf, color = created[0].get_volume()
    assert tsdf.shape == g[f"fu{case}_tsdf"].shape
    assert np.abs(tsdf - g[f"fu{case}_tsdf"]).max() <= 1e-5
    assert np.array_equal(created[0].get_weight_volume(), g[f"fu{case}_weight"])
    assert np.array_equal(color, g[f"fu{case}_color"])


def test_select_key_frames_matches_reference_golden():
    """hive_fusion.select_key_frames against the lists HiveDataset.select_key_frames itself returned."""
    from hive_b200 import hive_fusion
    from hive_b200.synthetic import _Trajectory
    g = _driver_gold()
    mats = g["kf_world_to_camera"]
    ds = SyntheticDataset(Scene("small", "room"), len(mats))
    ds.camera_trajectory = _Trajectory(list(mats))
    for a, thr in enumerate(g["kf_thresholds"]):
        for b, step in enumerate(g["kf_steps"]):
            want = [int(k) for k in g[f"kf_{a}_{b}"]]
            assert hive_fusion.select_key_frames(ds, float(thr), int(step)) == want, (thr, step)


def test_get_mesh_and_point_cloud_match_reference_golden(c_oracle):
    """TSDFVolume.get_mesh / get_point_cloud on exactly the planes the reference's driver produced, against what the
    reference's own get_mesh / get_point_cloud returned for them (fusion.py:311-350; the skimage call inside was answered
    by the C marching cubes, so topology parity is that of tests/test_marching_cubes_gpu.py -- what is pinned here is the
    post-processing: dtypes, world transform, colour gather and unpacking, the (V, 6) layout)."""
    import torch
    from hive_b200.fusion import TSDFVolume
    g = _driver_gold()
    tsdf, color = g["fu0_tsdf"], g["fu0_color"]
    vol = TSDFVolume(g["fu0_bounds"].copy(), float(g["fu0_voxel_size"]))
    X, Y, Z = tsdf.shape
    assert tuple(int(d) for d in vol._vol_dim) == (X, Y, Z)
    vol._tsdf_vol_gpu[:, :, :Z].copy_(torch.from_numpy(tsdf))
    vol._color_vol_gpu[:, :, :Z].copy_(torch.from_numpy(color))
    verts, faces, norms, colors = vol.get_mesh()
    # the golden is in the C oracle's vertex order, the GPU emits in (owner voxel, axis) order: renumber through the keys
    _, _, _, _, keys = c_oracle.marching_cubes(tsdf, 0.0, with_keys=True)
    order = c_oracle.product_order(keys)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    assert verts.dtype == np.float32 and faces.dtype == np.int32 and colors.dtype == np.uint8
    assert np.array_equal(verts, g["mp_verts"][order])
    assert np.array_equal(colors, g["mp_colors"][order])
    assert np.array_equal(faces, rank[g["mp_faces"]].astype(np.int32))
    assert np.abs(norms - g["mp_norms"][order]).max() <= 1e-6
    pc = vol.get_point_cloud()
    assert pc.dtype == g["mp_pc"].dtype and pc.shape == g["mp_pc"].shape
    assert np.array_equal(pc, g["mp_pc"][order])
