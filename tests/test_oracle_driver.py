"""CPU: pins the closed form of dilate_mask (oracle/driver_oracle.py) against OpenCV called exactly as the reference
calls it (hive/image_processing.py:41-42), and checks the driver restatement's voxel-budget rule (hive/fusion.py:66-74)."""
import numpy as np
import pytest

from oracle import driver_oracle as drv

cv2 = pytest.importorskip("cv2")


@pytest.mark.parametrize("shape", [(60, 80), (37, 45), (33, 64), (5, 3)])
@pytest.mark.parametrize("n", [0, 1, 2, 10, 40])
def test_closed_form_equals_cv2(shape, n):
    rng = np.random.default_rng(7 + n)
    mask = (rng.random(shape) < 0.02).astype(np.uint8) * rng.integers(1, 4, size=shape).astype(np.uint8)
    assert np.array_equal(drv.dilate_mask_numpy(mask, n), drv.dilate_mask_cv2(mask, n))


@pytest.mark.parametrize("n", [0, 3])
def test_edge_masks(n):
    for mask in (np.zeros((20, 31), np.uint8), np.ones((20, 31), np.uint8)):
        assert np.array_equal(drv.dilate_mask_numpy(mask, n), drv.dilate_mask_cv2(mask, n))
    corner = np.zeros((20, 70), np.uint8)
    corner[0, 0] = corner[19, 69] = 2
    got = drv.dilate_mask_numpy(corner, n)
    assert np.array_equal(got, drv.dilate_mask_cv2(corner, n))
    assert got.sum() == 2 * (n + 1) ** 2


def test_voxel_budget_rule():
    from hive_b200.synthetic import Scene, SyntheticDataset
    ds = SyntheticDataset(Scene("tiny", "room"), 3)
    vs, bounds = drv.adjust_voxel_size(ds, 0.01, None, range(3))
    assert vs == 0.01 and np.all(bounds[:, 0] <= 0) and np.all(bounds[:, 1] >= 0)  # always contains the origin
    vs2, bounds2 = drv.adjust_voxel_size(ds, 0.01, 100_000, range(3))
    assert np.array_equal(bounds, bounds2)
    assert vs2 > 0.01 and np.isclose(np.prod((bounds[:, 1] - bounds[:, 0]) / vs2), 100_000)


def test_srgb_table_is_the_reference_expression_over_its_domain():
    """hive/pipeline.py:281-282 is a function of one uint8; the table the CUDA path applies must equal the oracle's
    whole-array restatement for every input byte, in (V, 3) and (V, 4) layouts (alpha untouched)."""
    from hive_b200 import image_processing
    lut = image_processing.srgb_vertex_color_lut()
    assert lut.dtype == np.uint8 and lut.shape == (256,)
    assert lut[0] == 0 and lut[255] == 255 and np.all(np.diff(lut.astype(int)) >= 0)
    rng = np.random.default_rng(5)
    for cols in (3, 4):
        c = rng.integers(0, 256, (5000, cols), dtype=np.uint8)
        c[:256, 0] = np.arange(256)
        want = drv.vertex_colors_to_srgb(c)
        got = c.copy()
        got[:, :3] = lut[c[:, :3]]
        assert np.array_equal(got, want)
        assert want.dtype == np.uint8
