"""CPU: pins the numpy restatements of the Pipeline static methods (oracle/d2m_oracle.py) against the outputs of the
reference's own functions (tests/golden/pipeline_statics.npz, made by oracle/make_pipeline_golden.py from
hive/pipeline.py:651-694, 782-868 in the build container)."""
import os

import numpy as np
import pytest

from oracle import d2m_oracle as d2m

GOLD = os.path.join(os.path.dirname(__file__), "golden", "pipeline_statics.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def pack_case(gold, i):
    n = int(gold[f"pk{i}_n"])
    return ([gold[f"pk{i}_tex{k}"] for k in range(n)], [gold[f"pk{i}_uv{k}"] for k in range(n)], int(gold[f"pk{i}_rows"]),
            gold[f"pk{i}_atlas"], gold[f"pk{i}_final"])


def test_filter_faces_restatement(gold):
    kept = d2m.filter_faces(gold["ff_points2d"], gold["ff_depth"], gold["ff_raw"])
    assert np.array_equal(kept, gold["ff_kept"])
    assert 0 < len(kept) < len(gold["ff_raw"])


def test_texture_and_uv_restatement(gold):
    from hive_b200.synthetic import Scene
    sc = Scene("small", "room")
    texture, uv = d2m.get_mesh_texture_and_uv(gold["tx_vertices"], sc.frame(2).rgb, sc.K, gold["tx_R"], gold["tx_t"])
    assert np.array_equal(texture, gold["tx_texture"])
    assert uv.dtype == gold["tx_uv"].dtype and np.array_equal(uv, gold["tx_uv"])


def test_pack_textures_restatement(gold):
    for i in range(int(gold["pk_cases"])):
        tex, uvs, n_rows, atlas, final = pack_case(gold, i)
        before = [u.copy() for u in uvs]
        got_atlas, got_final = d2m.pack_textures(tex, uvs, n_rows)
        assert got_atlas.dtype == np.uint8 and np.array_equal(got_atlas, atlas), i
        assert got_final.dtype == np.float64 and np.array_equal(got_final, final), i
        assert all(np.array_equal(a, b) for a, b in zip(uvs, before))


def test_pack_textures_empty_row_is_an_index_error():
    tex = [np.ones((2, 2, 3), np.uint8)] * 2
    uvs = [np.zeros((1, 2), np.int32)] * 2
    with pytest.raises(IndexError):
        d2m.pack_textures(tex, uvs, n_rows=2)  # 2 slots per row: the second row receives nothing
    with pytest.raises(IndexError):
        d2m.pack_textures([], [], n_rows=1)


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree absent")
def test_reference_raises_index_error_too():
    from oracle import ref_shim
    ref = ref_shim.reference_pipeline_statics(("_pack_textures",))["_pack_textures"]
    tex = [np.ones((2, 2, 3), np.uint8)] * 2
    with pytest.raises(IndexError):
        ref(tex, [np.zeros((1, 2), np.int32), np.zeros((1, 2), np.int32)], n_rows=2)
    with pytest.raises(IndexError):
        ref([], [], n_rows=1)
