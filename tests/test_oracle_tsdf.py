"""Pin the TSDF oracles (oracle/fusion_oracle.py numpy, oracle/tsdf_oracle.c) against the reference.

* golden vectors under tests/golden/ were produced by the reference's own CPU path (oracle/make_golden.py);
* when /root/reference is present (build container) the reference is additionally run live.
"""
import hashlib
import os

import numpy as np
import pytest

from hive_b200.synthetic import Scene
from oracle import fusion_oracle as fo
from oracle import ref_shim

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _run(vol, variant, weights):
    scene = Scene("tiny", variant)
    for i in range(scene.num_frames):
        fr = scene.frame(i)
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose, weights[i % len(weights)])
    return vol


@pytest.mark.parametrize("variant", ["room", "dense"])
@pytest.mark.parametrize("impl", ["numpy", "c", "numba"])
def test_oracle_matches_reference_golden(variant, impl, c_oracle):
    g = np.load(os.path.join(GOLDEN, f"tsdf_tiny_{variant}.npz"))
    assert _sha(g["tsdf"]) == str(g["sha_tsdf"]) and _sha(g["color"]) == str(g["sha_color"])
    scene = Scene("tiny", variant)
    if impl == "numba":
        fnb = pytest.importorskip("oracle.fusion_numba")  # the reference-shaped restatement bench.py --impl reference times
        cls = fnb.NumbaTSDFVolume
    else:
        cls = fo.OracleTSDFVolume if impl == "numpy" else c_oracle.COracleTSDFVolume
    vol = _run(cls(scene.vol_bnds.copy(), scene.voxel_size), variant, list(g["weights"]))
    assert tuple(vol._vol_dim) == tuple(g["vol_dim"])
    assert np.array_equal(vol._vol_origin, g["vol_origin"])
    # the restatement is bit-exact with the reference on all three planes
    assert np.array_equal(vol._weight_vol_cpu, g["weight"])
    assert np.array_equal(vol._color_vol_cpu, g["color"])
    assert np.array_equal(vol._tsdf_vol_cpu, g["tsdf"])
    assert (g["weight"] > 0).sum() > 1000  # the fixture is not trivial


def test_numpy_and_c_oracles_agree_on_a_larger_volume(c_oracle):
    scene = Scene("small", "room")
    a = fo.OracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    b = c_oracle.COracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for i in range(3):
        fr = scene.frame(i)
        a.integrate(fr.rgb, fr.depth, scene.K, fr.pose, 1.0)
        b.integrate(fr.rgb, fr.depth, scene.K, fr.pose, 1.0)
        assert a.last_updated == b.last_updated > 0
    assert np.array_equal(a._tsdf_vol_cpu, b._tsdf_vol_cpu)
    assert np.array_equal(a._weight_vol_cpu, b._weight_vol_cpu)
    assert np.array_equal(a._color_vol_cpu, b._color_vol_cpu)


def test_frustum_matches_reference_golden():
    g = np.load(os.path.join(GOLDEN, "frustum.npz"))
    scene = Scene("tiny", "room")
    from hive_b200 import fusion as product_host  # host-only helper, needs no GPU
    for i in range(3):
        fr = scene.frame(i)
        assert np.array_equal(fo.get_view_frustum(fr.depth, scene.K, fr.pose), g[f"f{i}"])
        assert np.array_equal(product_host.get_view_frustum(fr.depth, scene.K, fr.pose), g[f"f{i}"])


def test_volume_geometry_known_answer():
    """The one known answer in the reference tree: 405 x 264 x 289 voxels at 2 cm
    (third_party/tsdf_fusion_python/README.md:22) from the bounds of the shipped 7-Scenes frames."""
    bnds = np.array([[-4.2211, 3.8680], [-2.6663, 2.6015], [0.0, 5.7627]])
    _, dim, origin = fo.volume_geometry(bnds, 0.02)
    assert tuple(dim) == (405, 264, 289)
    assert origin.dtype == np.float32


@pytest.mark.skipif(not ref_shim.available(), reason="/root/reference is only present in the build container")
def test_oracle_matches_live_reference(c_oracle):
    fusion = ref_shim.reference_fusion()
    scene = Scene("tiny", "room")
    ref = fusion.TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, use_gpu=False)
    orc = c_oracle.COracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for i in range(3):
        fr = scene.frame(i)
        w = 1.0 if i != 1 else 0.7  # a non-integer observation weight too
        ref.integrate(fr.rgb, fr.depth, scene.K, fr.pose, obs_weight=w)
        orc.integrate(fr.rgb, fr.depth, scene.K, fr.pose, w)
    assert np.array_equal(ref._weight_vol_cpu, orc._weight_vol_cpu)
    assert np.array_equal(ref._color_vol_cpu, orc._color_vol_cpu)
    assert np.array_equal(ref._tsdf_vol_cpu, orc._tsdf_vol_cpu)


@pytest.mark.parametrize("impl", ["numpy", "c", "numba"])
def test_oracle_matches_reference_on_7scenes_fixture(impl, c_oracle):
    """Real sensor data: the reference's own 7-Scenes frames (third_party/tsdf_fusion_python/data, read as
    demo.py:19-55 does, subsampled 4x) through the reference's CPU path -> tests/golden/tsdf_7scenes.npz."""
    g = np.load(os.path.join(GOLDEN, "tsdf_7scenes.npz"))
    assert _sha(g["tsdf"]) == str(g["sha_tsdf"]) and _sha(g["color"]) == str(g["sha_color"])
    if impl == "numba":
        fnb = pytest.importorskip("oracle.fusion_numba")  # the reference-shaped restatement bench.py --impl reference times
        cls = fnb.NumbaTSDFVolume
    else:
        cls = fo.OracleTSDFVolume if impl == "numpy" else c_oracle.COracleTSDFVolume
    vol = cls(g["vol_bnds"].copy(), float(g["voxel_size"]))
    for d, c, p in zip(g["depth"], g["rgb"], g["pose"]):
        vol.integrate(c, d, g["cam_intr"], p, 1.0)
    assert tuple(vol._vol_dim) == tuple(g["vol_dim"]) and np.array_equal(vol._vol_origin, g["vol_origin"])
    assert np.array_equal(vol._weight_vol_cpu, g["weight"])
    assert np.array_equal(vol._color_vol_cpu, g["color"])
    assert np.array_equal(vol._tsdf_vol_cpu, g["tsdf"])
    assert (g["weight"] > 0).sum() > 30_000
