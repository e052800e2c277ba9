"""Golden vectors of the reference's own ``Pipeline`` static methods (build container only).

    python -m oracle.make_pipeline_golden

TEST INFRASTRUCTURE.  ``hive/pipeline.py`` cannot be imported here (trimesh, openmesh, detectron2), so
``oracle/ref_shim.reference_pipeline_statics`` compiles the four static methods on the path straight from the reference's
source file and this script runs them on seeded inputs.  Output: ``tests/golden/pipeline_statics.npz``

``ff_*``    ``_filter_faces(points2d, depth, _triangulate_faces(points2d), options)`` (pipeline.py:651-694) on one frame
            of the ``small``/``room`` scene, default options (2 px, 0.1 m) -- inputs, the raw Delaunay faces and the
            kept faces (the raw faces carry this scipy's diagonal choice; the kept set given those faces is exact)
``tx_*``    ``_get_mesh_texture_and_uv`` (pipeline.py:782-808) of that frame's point cloud over its colour image
``pk<i>_*`` ``_pack_textures`` (pipeline.py:810-868): three crops in one row, a single crop, five crops in two rows,
            float64 UVs, crops without vertices / of zero width
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from hive_b200.synthetic import Scene  # noqa: E402
from oracle import ref_shim  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


class _Options:
    max_pixel_distance = 2
    max_depth_distance = 0.1


def pack_cases():
    rng = np.random.default_rng(20261104)

    def crops(shapes, n_uv, dtype=np.int32):
        tex = [rng.integers(1, 256, (h, w, 3), dtype=np.uint8) for h, w in shapes]
        uvs = []
        for (h, w), n in zip(shapes, n_uv):
            uv = np.stack([rng.integers(0, max(w, 1), n), rng.integers(0, max(h, 1), n)], axis=1)
            uvs.append(uv.astype(dtype) + (0.25 if dtype == np.float64 else 0))
        return tex, uvs

    yield crops([(7, 5), (3, 9), (12, 4)], [11, 4, 30]), 1
    yield crops([(16, 16)], [40]), 1
    yield crops([(7, 5), (3, 9), (12, 4), (2, 2), (9, 31)], [5, 6, 7, 8, 9]), 2
    yield crops([(33, 17), (5, 70)], [100, 3], np.float64), 1
    yield crops([(4, 4), (6, 0), (5, 3), (2, 8)], [3, 0, 0, 6]), 1


def main():
    assert ref_shim.available(), "the reference tree is needed to generate golden vectors"
    ref = ref_shim.reference_pipeline_statics()
    geo = ref_shim.reference_geometric()
    out = {}

    scene = Scene("small", "room")
    f = scene.frame(2)
    w2c = np.linalg.inv(f.pose)
    R, t = w2c[:3, :3], w2c[:3, 3:]
    mask = f.mask > 0
    valid = mask & (f.depth > 0)
    V, U = valid.nonzero()
    points2d = np.array([U, V]).T  # pipeline.py:393-397
    depth = f.depth[valid]
    raw = ref["_triangulate_faces"](points2d)
    kept = ref["_filter_faces"](points2d, depth, raw, _Options())
    out.update(ff_frame=2, ff_points2d=points2d.astype(np.int32), ff_depth=depth, ff_raw=raw.astype(np.int32),
               ff_kept=np.ascontiguousarray(kept).astype(np.int32))

    verts = geo.point_cloud_from_depth(f.depth, mask, scene.K, R, t)
    texture, uv = ref["_get_mesh_texture_and_uv"](verts, f.rgb, scene.K, R, t)
    out.update(tx_vertices=verts, tx_R=R, tx_t=t, tx_texture=texture, tx_uv=uv)

    n_cases = 0
    for i, ((tex, uvs), n_rows) in enumerate(pack_cases()):
        for k, (a, u) in enumerate(zip(tex, uvs)):
            out[f"pk{i}_tex{k}"] = a
            out[f"pk{i}_uv{k}"] = u.copy()
        atlas, final = ref["_pack_textures"](tex, [u.copy() for u in uvs], n_rows=n_rows)
        out[f"pk{i}_n"] = len(tex)
        out[f"pk{i}_rows"] = n_rows
        out[f"pk{i}_atlas"] = atlas
        out[f"pk{i}_final"] = final
        n_cases += 1
    out["pk_cases"] = n_cases

    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, "pipeline_statics.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", len(raw), "raw faces,", len(kept), "kept; texture", texture.shape)


if __name__ == "__main__":
    main()
