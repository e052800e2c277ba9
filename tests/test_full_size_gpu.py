"""BASELINE config C2 at full size (512^3 voxels, 640x480 frames): too large for the oracle to finish in seconds, so the
CUDA path is checked through size-independent properties of the domain (task statement, parity rule 3):

* integrate: all kernel variants agree (weight/colour bit-identical), integrating the same observation twice leaves
  the running mean where it was and doubles the weight, untouched voxels keep their initial state, the update counter
  equals the number of voxels whose weight changed;
* marching cubes: one vertex per sign-changing lattice edge (counted independently with numpy), every triangle edge is
  used by at most two triangles and by exactly two away from the volume border, vertex colours equal the nearest
  voxel's unpacked colour.
"""
import numpy as np
import pytest

from hive_b200.synthetic import Scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2_frames():
    scene = Scene("C2", "room")
    return scene, [scene.frame(i) for i in (0, 7)]


def test_c2_integrate_properties(c2_frames):
    from hive_b200.fusion import TSDFVolume
    scene, frames = c2_frames
    planes = {}
    for kernel in (1, 3):
        vol = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
        assert tuple(vol._vol_dim) == (512, 512, 512)
        vol.integrate_kernel = kernel
        vol.count_updates = True
        fr = frames[0]
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        t1, c1 = vol.get_volume()
        w1 = vol.get_weight_volume()
        n1 = vol.num_updated()
        assert n1 == int((w1 > 0).sum()) > 5_000_000
        untouched = w1 == 0
        assert (t1[untouched] == 1.0).all() and (c1[untouched] == 0.0).all()
        assert (w1[~untouched] == 1.0).all() and (t1 <= 1.0).all() and (t1 >= -1.0).all()
        # the same observation again: the mean stays, the weight doubles
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        t2, c2 = vol.get_volume()
        w2 = vol.get_weight_volume()
        assert np.array_equal(w2, 2.0 * w1) and np.array_equal(c2, c1)
        assert np.abs(t2 - t1).max() <= 1e-6
        assert vol.num_updated() == 2 * n1
        fr = frames[1]
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose, obs_weight=3.0)
        planes[kernel] = (vol.get_volume(), vol.get_weight_volume())
        del vol
    (ta, ca), wa = planes[1]
    (tb, cb), wb = planes[3]
    assert np.array_equal(wa, wb) and np.array_equal(ca, cb)
    assert np.abs(ta - tb).max() <= 1e-5


def test_c2_mesh_properties(c2_frames):
    from hive_b200.fusion import TSDFVolume
    scene, frames = c2_frames
    vol = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for fr in frames:
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
    verts, faces, norms, colors = vol.get_mesh()
    tsdf, color = vol.get_volume()
    pos = tsdf > 0
    expected = int((pos[1:] != pos[:-1]).sum() + (pos[:, 1:] != pos[:, :-1]).sum() + (pos[:, :, 1:] != pos[:, :, :-1]).sum())
    assert len(verts) >= expected  # + the few cube-centre vertices of the 12-gon tilings
    assert len(verts) - expected < 0.001 * expected + 16
    assert faces.min() == 0 and faces.max() == len(verts) - 1 and len(faces) > 500_000
    # manifoldness: directed edges are unique; an undirected edge without a partner lies on the volume border
    e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]]).astype(np.int64)
    fwd = e[:, 0] * len(verts) + e[:, 1]
    rev = e[:, 1] * len(verts) + e[:, 0]
    assert len(np.unique(fwd)) == len(fwd)
    lonely = e[~np.isin(fwd, rev)]
    if len(lonely):
        vox = (verts[lonely.reshape(-1)] - vol._vol_origin) / vol._voxel_size
        border = (vox < 1.0).any(axis=1) | (vox > np.array(tsdf.shape) - 2.0).any(axis=1)
        assert border.all()
    # colours: nearest voxel, b*65536 + g*256 + r unpacked (fusion.py:339-348)
    idx = np.round((verts.astype(np.float64) - vol._vol_origin) / vol._voxel_size).astype(int)
    idx = np.clip(idx, 0, np.array(tsdf.shape) - 1)
    packed = color[idx[:, 0], idx[:, 1], idx[:, 2]]
    b = np.floor(packed / 65536.0)
    g = np.floor((packed - b * 65536.0) / 256.0)
    r = packed - b * 65536.0 - g * 256.0
    match = (colors == np.stack([r, g, b], axis=1).astype(np.uint8)).all(axis=1)
    assert match.mean() > 0.999  # (the world -> voxel round trip above is float64; the product rounds the float32 voxel coordinate)
    ln = np.linalg.norm(norms, axis=1)
    assert np.isfinite(norms).all() and ln.max() < 1.0 + 1e-3 and (np.abs(ln - 1.0) < 1e-3).mean() > 0.999
