"""On-disk formats: the glb writer round-trips through its own reader and obeys the glTF 2.0 binary
container rules (CPU); the plane checkpoint resumes an integration bit for bit (GPU)."""
import json
import struct

import numpy as np
import pytest

from hive_b200 import export


def _mesh(n=37, seed=0):
    rng = np.random.default_rng(seed)
    verts = rng.standard_normal((n, 3)).astype(np.float32)
    faces = rng.integers(0, n, size=(2 * n + 1, 3)).astype(np.int32)
    colors = rng.integers(0, 256, size=(n, 3)).astype(np.uint8)
    normals = rng.standard_normal((n, 3)).astype(np.float32)
    return verts, faces, colors, normals


def test_glb_container_and_round_trip(tmp_path):
    verts, faces, colors, normals = _mesh()
    path = tmp_path / "mesh.glb"
    n = export.write_glb(path, verts.astype(np.float64), faces, colors, normals)
    data = path.read_bytes()
    assert len(data) == n and n % 4 == 0
    magic, version, total = struct.unpack_from("<III", data, 0)
    assert (magic, version, total) == (0x46546C67, 2, n)
    n_json, kind = struct.unpack_from("<II", data, 12)
    assert kind == 0x4E4F534A and n_json % 4 == 0
    doc = json.loads(data[20:20 + n_json])
    assert doc["asset"]["version"] == "2.0" and doc["meshes"][0]["primitives"][0]["mode"] == 4
    for view in doc["bufferViews"]:
        assert view["byteOffset"] % 4 == 0 and view["byteOffset"] + view["byteLength"] <= doc["buffers"][0]["byteLength"]
    pos = doc["accessors"][doc["meshes"][0]["primitives"][0]["attributes"]["POSITION"]]
    assert np.allclose(pos["min"], verts.min(axis=0)) and np.allclose(pos["max"], verts.max(axis=0))  # required by the spec
    _, arrays, image = export.read_glb(path)
    assert image is None
    assert np.array_equal(arrays["POSITION"], verts) and np.array_equal(arrays["NORMAL"], normals)
    assert np.array_equal(arrays["COLOR_0"][:, :3], colors) and np.all(arrays["COLOR_0"][:, 3] == 255)
    assert np.array_equal(arrays["indices"].reshape(-1, 3), faces)


def test_glb_textured_mesh_and_odd_sizes(tmp_path):
    verts, faces, _, _ = _mesh(n=5, seed=3)
    uv = np.linspace(0, 1, 10).reshape(5, 2)
    png = bytes(range(1, 8))  # 7 bytes: the image view must still leave the buffer 4-byte aligned
    path = tmp_path / "fg.glb"
    export.write_glb(path, verts, faces[:1], uv=uv, texture_png=png, name="frame_000")
    doc, arrays, image = export.read_glb(path)
    assert image == png and doc["materials"][0]["pbrMetallicRoughness"]["baseColorTexture"]["index"] == 0
    assert np.allclose(arrays["TEXCOORD_0"], uv) and arrays["indices"].shape == (3,)
    with pytest.raises(AssertionError):
        export.write_glb(path, verts, np.array([[0, 1, 5]]))


@pytest.mark.gpu
def test_checkpoint_resumes_bit_for_bit(tmp_path):
    from hive_b200.fusion import TSDFVolume
    from hive_b200.synthetic import Scene
    scene = Scene("small", "room")
    frames = [scene.frame(i) for i in range(6)]
    for z_range in (None, (8, 64)):
        whole = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, z_range=z_range)
        first = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, z_range=z_range)
        for fr in frames:
            whole.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        for fr in frames[:3]:
            first.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        path = tmp_path / f"ckpt_{0 if z_range is None else 1}.npz"
        export.save_checkpoint(first, path)
        del first
        resumed = export.load_checkpoint(path)
        for fr in frames[3:]:
            resumed.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        a, b = whole.get_volume(), resumed.get_volume()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.array_equal(whole.get_weight_volume(), resumed.get_weight_volume())
        if z_range is None:
            va, fa, na, ca = whole.get_mesh()
            vb, fb, nb, cb = resumed.get_mesh()
            assert np.array_equal(va, vb) and np.array_equal(fa, fb) and np.array_equal(ca, cb)
            # and the mesh goes out as glb: positions, colours and triangles come back as they went in
            glb = tmp_path / "background.glb"
            export.write_glb(glb, vb, fb, cb, nb)
            _, arrays, _ = export.read_glb(glb)
            assert np.array_equal(arrays["POSITION"], vb) and np.array_equal(arrays["indices"].reshape(-1, 3), fb)
            assert np.array_equal(arrays["COLOR_0"][:, :3], cb)
