"""meshwrite / pcwrite (fusion.py:378-438): the C writers produce byte-identical ASCII files to the reference's Python
loops -- against the restatement in oracle/fusion_oracle.py and, in the build container, against the reference's own
functions.  Host-only code: runs without a GPU."""
import struct

import numpy as np
import pytest

from oracle import fusion_oracle as fo
from oracle import ref_shim


def _mesh(n_verts, n_faces, seed=0):
    rng = np.random.default_rng(seed)
    verts = (rng.standard_normal((n_verts, 3)) * 3).astype(np.float32)
    verts[:7] = [[0, -0.0, 1e-7], [123456.789, -98765.4321, 0.5], [1e10, -1e10, 2.5e-6], [0.0000005, 0.0000015, 0.0000025],
                 [-0.0000005, 1.9999995, 3.4e38], [1 / 3, 2 / 3, -1 / 3], [0.125, 0.375, 0.4999995]]
    norms = rng.standard_normal((n_verts, 3)).astype(np.float32)
    norms /= np.linalg.norm(norms, axis=1, keepdims=True)
    colors = rng.integers(0, 256, size=(n_verts, 3)).astype(np.uint8)
    faces = rng.integers(0, n_verts, size=(n_faces, 3)).astype(np.int32)
    return verts, faces, norms, colors


@pytest.mark.parametrize("n_verts,n_faces", [(7, 0), (50, 80), (20_000, 39_000)])
def test_meshwrite_is_byte_identical(n_verts, n_faces, tmp_path, hv_lib):
    from hive_b200 import fusion
    verts, faces, norms, colors = _mesh(n_verts, n_faces)
    fusion.meshwrite(str(tmp_path / "a.ply"), verts, faces, norms, colors)
    fo.meshwrite(str(tmp_path / "b.ply"), verts, faces, norms, colors)
    assert (tmp_path / "a.ply").read_bytes() == (tmp_path / "b.ply").read_bytes()


@pytest.mark.parametrize("n", [1, 33, 20_000])
def test_pcwrite_is_byte_identical(n, tmp_path, hv_lib):
    from hive_b200 import fusion
    verts, _, _, colors = _mesh(max(n, 7), 0, seed=n)
    xyzrgb = np.hstack([verts[:n], colors[:n]])  # what get_point_cloud returns (fusion.py:329)
    fusion.pcwrite(str(tmp_path / "a.ply"), xyzrgb)
    fo.pcwrite(str(tmp_path / "b.ply"), xyzrgb)
    assert (tmp_path / "a.ply").read_bytes() == (tmp_path / "b.ply").read_bytes()


@pytest.mark.skipif(not ref_shim.available(), reason="/root/reference is only present in the build container")
def test_writers_match_the_live_reference(tmp_path, hv_lib):
    from hive_b200 import fusion
    ref = ref_shim.reference_fusion()
    verts, faces, norms, colors = _mesh(3000, 5000, seed=5)
    fusion.meshwrite(str(tmp_path / "a.ply"), verts, faces, norms, colors)
    ref.meshwrite(str(tmp_path / "r.ply"), verts, faces, norms, colors)
    assert (tmp_path / "a.ply").read_bytes() == (tmp_path / "r.ply").read_bytes()
    xyzrgb = np.hstack([verts, colors])
    fusion.pcwrite(str(tmp_path / "pa.ply"), xyzrgb)
    ref.pcwrite(str(tmp_path / "pr.ply"), xyzrgb)
    assert (tmp_path / "pa.ply").read_bytes() == (tmp_path / "pr.ply").read_bytes()


def test_binary_ply_round_trip(tmp_path, hv_lib):
    from hive_b200 import fusion
    verts, faces, norms, colors = _mesh(100, 150, seed=2)
    path = tmp_path / "m.ply"
    fusion.meshwrite(str(path), verts, faces, norms, colors, binary=True)
    raw = path.read_bytes()
    head, body = raw.split(b"end_header\n", 1)
    assert b"format binary_little_endian 1.0" in head and b"element vertex 100" in head and b"element face 150" in head
    assert len(body) == 100 * 27 + 150 * 13
    rec = struct.unpack_from("<6f3B", body, 27 * 42)
    assert np.array_equal(np.float32(rec[:3]), verts[42]) and tuple(rec[6:]) == tuple(colors[42])
    cnt, a, b, c = struct.unpack_from("<B3i", body, 100 * 27 + 13 * 17)
    assert cnt == 3 and (a, b, c) == tuple(faces[17])


def test_unwritable_path_raises(tmp_path, hv_lib):
    from hive_b200 import fusion
    verts, faces, norms, colors = _mesh(8, 2)
    with pytest.raises(AssertionError):
        fusion.meshwrite(str(tmp_path / "no_such_dir" / "m.ply"), verts, faces, norms, colors)
