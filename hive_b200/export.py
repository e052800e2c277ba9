"""On-disk formats next to the hot path: a binary glTF (``.glb``) writer for vertex-coloured or textured
triangle meshes and a plane checkpoint of a ``TSDFVolume``.

The reference writes its meshes as ``.glb`` through ``trimesh.exchange.export.export_scene`` (hive/pipeline.py:920-934);
trimesh is a third-party dependency that is installed on neither box, so this is a writer of the *format* (glTF 2.0, one
binary buffer), not a byte-for-byte copy of trimesh's exporter -- PARITY UNPINNED.  What the tests pin: the file parses as
glTF 2.0 (header, chunk lengths, 4-byte alignment, accessor min/max) and the arrays read back equal the arrays written.
Host code: a few hundred bytes of JSON and one ``tobytes`` per array.

The plane checkpoint has no counterpart in the reference (its volume dies with the process); it stores exactly the state
``integrate`` carries from frame to frame.
"""
from __future__ import annotations

import json
import struct

import numpy as np

_GLB_MAGIC, _JSON_CHUNK, _BIN_CHUNK = 0x46546C67, 0x4E4F534A, 0x004E4942
_FLOAT, _UBYTE, _UINT = 5126, 5121, 5125
_ARRAY_BUFFER, _ELEMENT_ARRAY_BUFFER = 34962, 34963


def _pad4(b: bytes, fill: bytes) -> bytes:
    return b + fill * ((-len(b)) % 4)


def write_glb(filename, verts, faces, colors=None, normals=None, uv=None, texture_png: bytes = None, name="mesh"):
    """One triangle mesh as a binary glTF 2.0 file.

    ``verts`` (V, 3) float, ``faces`` (F, 3) int, optional ``colors`` (V, 3|4) uint8 (COLOR_0, normalised), ``normals``
    (V, 3) float, ``uv`` (V, 2) float with ``texture_png`` (the encoded image bytes of the atlas: encoding images is not
    this package's business) for a textured foreground mesh.  Returns the number of bytes written.
    """
    verts = np.ascontiguousarray(verts, dtype="<f4")
    faces = np.ascontiguousarray(faces, dtype="<u4")
    assert verts.ndim == 2 and verts.shape[1] == 3 and faces.ndim == 2 and faces.shape[1] == 3
    n_verts = int(verts.shape[0])
    assert faces.size == 0 or int(faces.max()) < n_verts, "a face refers to a vertex that does not exist"
    blob, views, accessors = bytearray(), [], []

    def add(array, component, kind, target, normalized=False, minmax=False):
        raw = _pad4(array.tobytes(), b"\x00")
        views.append({"buffer": 0, "byteOffset": len(blob), "byteLength": array.nbytes, "target": target})
        acc = {"bufferView": len(views) - 1, "componentType": component, "count": int(array.shape[0]), "type": kind}
        if normalized:
            acc["normalized"] = True
        if minmax and array.shape[0]:
            acc["min"] = [float(v) for v in array.min(axis=0)]
            acc["max"] = [float(v) for v in array.max(axis=0)]
        accessors.append(acc)
        blob.extend(raw)
        return len(accessors) - 1

    attributes = {"POSITION": add(verts, _FLOAT, "VEC3", _ARRAY_BUFFER, minmax=True)}
    if normals is not None:
        normals = np.ascontiguousarray(normals, dtype="<f4")
        assert normals.shape == verts.shape
        attributes["NORMAL"] = add(normals, _FLOAT, "VEC3", _ARRAY_BUFFER)
    if colors is not None:
        colors = np.asarray(colors)
        assert colors.shape[0] == n_verts and colors.shape[1] in (3, 4)
        rgba = np.full((n_verts, 4), 255, dtype=np.uint8)  # 4 components: rows stay 4-byte aligned
        rgba[:, :colors.shape[1]] = colors
        attributes["COLOR_0"] = add(rgba, _UBYTE, "VEC4", _ARRAY_BUFFER, normalized=True)
    if uv is not None:
        uv = np.ascontiguousarray(uv, dtype="<f4")
        assert uv.shape == (n_verts, 2)
        attributes["TEXCOORD_0"] = add(uv, _FLOAT, "VEC2", _ARRAY_BUFFER)
    indices = add(faces.reshape(-1), _UINT, "SCALAR", _ELEMENT_ARRAY_BUFFER)

    primitive = {"attributes": attributes, "indices": indices, "mode": 4}
    doc = {"asset": {"version": "2.0", "generator": "hive_b200"}, "scene": 0, "scenes": [{"nodes": [0]}],
           "nodes": [{"mesh": 0, "name": name}], "meshes": [{"name": name, "primitives": [primitive]}],
           "accessors": accessors, "bufferViews": views}
    if texture_png is not None:
        assert uv is not None, "a texture needs uv coordinates"
        views.append({"buffer": 0, "byteOffset": len(blob), "byteLength": len(texture_png)})
        blob.extend(_pad4(bytes(texture_png), b"\x00"))
        doc["images"] = [{"bufferView": len(views) - 1, "mimeType": "image/png"}]
        doc["samplers"] = [{"magFilter": 9729, "minFilter": 9729}]
        doc["textures"] = [{"source": 0, "sampler": 0}]
        doc["materials"] = [{"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}, "metallicFactor": 0.0,
                                                      "roughnessFactor": 1.0}}]
        primitive["material"] = 0
    doc["buffers"] = [{"byteLength": len(blob)}]
    js = _pad4(json.dumps(doc, separators=(",", ":")).encode("utf-8"), b" ")
    total = 12 + 8 + len(js) + 8 + len(blob)
    with open(filename, "wb") as f:
        f.write(struct.pack("<III", _GLB_MAGIC, 2, total))
        f.write(struct.pack("<II", len(js), _JSON_CHUNK))
        f.write(js)
        f.write(struct.pack("<II", len(blob), _BIN_CHUNK))
        f.write(bytes(blob))
    return total


def read_glb(filename):
    """Parse what ``write_glb`` writes (tests; a viewer would do the same): returns (json document, {attribute or
    'indices': array}, image bytes or None)."""
    with open(filename, "rb") as f:
        data = f.read()
    magic, version, total = struct.unpack_from("<III", data, 0)
    assert magic == _GLB_MAGIC and version == 2 and total == len(data), "not a glTF 2.0 binary file"
    n_json, kind = struct.unpack_from("<II", data, 12)
    assert kind == _JSON_CHUNK and n_json % 4 == 0
    doc = json.loads(data[20:20 + n_json].decode("utf-8"))
    n_bin, kind = struct.unpack_from("<II", data, 20 + n_json)
    assert kind == _BIN_CHUNK and n_bin % 4 == 0 and 28 + n_json + n_bin == total
    blob = data[28 + n_json:]
    assert doc["buffers"][0]["byteLength"] == n_bin
    dtypes = {_FLOAT: "<f4", _UBYTE: "u1", _UINT: "<u4"}
    width = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4}

    def load(index):
        acc = doc["accessors"][index]
        view = doc["bufferViews"][acc["bufferView"]]
        assert view["byteOffset"] % 4 == 0
        a = np.frombuffer(blob, dtype=dtypes[acc["componentType"]], count=acc["count"] * width[acc["type"]],
                          offset=view["byteOffset"])
        return a.reshape(acc["count"], width[acc["type"]]) if width[acc["type"]] > 1 else a

    prim = doc["meshes"][0]["primitives"][0]
    arrays = {k: load(v) for k, v in prim["attributes"].items()}
    arrays["indices"] = load(prim["indices"])
    image = None
    if "images" in doc:
        view = doc["bufferViews"][doc["images"][0]["bufferView"]]
        image = bytes(blob[view["byteOffset"]:view["byteOffset"] + view["byteLength"]])
    return doc, arrays, image


# ---- plane checkpoint -----------------------------------------------------------------------------------------------------
_CKPT_VERSION = 1


def save_checkpoint(vol, filename):
    """Everything ``TSDFVolume.integrate`` carries from frame to frame -- the three planes of this volume (or z-slab) and
    the parameters they were built with -- as one ``.npz`` (uncompressed: the planes are the bulk and do not compress)."""
    tsdf, color = vol.get_volume()
    np.savez(filename, version=np.int64(_CKPT_VERSION), tsdf=tsdf, weight=vol.get_weight_volume(), color=color,
             vol_bnds=np.asarray(vol._vol_bnds, dtype=np.float64), vol_dim=np.asarray(vol._vol_dim, dtype=np.int64),
             voxel_size=np.float64(vol._voxel_size),
             z_range=np.array([vol._z0, vol._z1], dtype=np.int64),
             weights=np.array([1.0 if vol._integer_weights else 0.0, vol._weight_sum], dtype=np.float64))


def load_checkpoint(filename, device=None):
    """A ``TSDFVolume`` in the state ``save_checkpoint`` saw: integrating the remaining frames into it gives, bit for bit,
    what the uninterrupted volume gives."""
    from . import _xfer
    from .fusion import TSDFVolume
    with np.load(filename) as z:
        assert int(z["version"]) == _CKPT_VERSION, "unknown checkpoint version"
        vs = float(z["voxel_size"])
        # upper bounds half a voxel inside the last one: the constructor's ceil() then gives the saved dimensions whatever
        # the rounding of (dim * vs) / vs, and it recomputes the upper bounds from them (fusion.py:41-43)
        bnds = z["vol_bnds"].copy()
        bnds[:, 1] = bnds[:, 0] + (z["vol_dim"] - 0.5) * vs
        z0, z1 = (int(v) for v in z["z_range"])
        vol = TSDFVolume(bnds, vs, device=device, z_range=None if (z0, z1) == (0, int(z["vol_dim"][2])) else (z0, z1))
        assert np.array_equal(vol._vol_dim, z["vol_dim"]) and np.array_equal(vol._vol_origin, z["vol_bnds"][:, 0].astype(np.float32))
        Z = int(vol._local_dim[2])
        for name, plane in (("tsdf", vol._tsdf_vol_gpu), ("weight", vol._weight_vol_gpu), ("color", vol._color_vol_gpu)):
            host = z[name]
            assert host.shape == tuple(int(d) for d in vol._local_dim), f"checkpoint plane `{name}` does not fit the volume"
            plane[:, :, :Z].copy_(_xfer.to_device(np.ascontiguousarray(host, dtype=np.float32), vol.device))
        vol._integer_weights, vol._weight_sum = bool(z["weights"][0]), float(z["weights"][1])
    return vol
