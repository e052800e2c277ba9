"""ORACLE -- test infrastructure, not product code.

CPU restatement of HIVE's per-frame depth-to-mesh step:

* ``point_cloud_from_depth`` / ``image2world`` / ``world2image``  -- hive/geometric.py:107-126, 183-206, 155-180
  (verbatim numpy; pinned against the real module through tests/golden/backproject.npz, oracle/make_golden.py);
* ``triangulate_faces`` / ``filter_faces``  -- Pipeline._triangulate_faces / _filter_faces, hive/pipeline.py:651-694
  (``hive.pipeline`` itself cannot be imported: trimesh, openmesh, detectron2 are absent; both are 10-line static
  methods around ``scipy.spatial.Delaunay``).  PARITY UNPINNED for the triangulation: Qhull lives in scipy
  (pinned 1.7.3 in requirements.txt:2, 1.18 here), the reference has no test for it, and on the pixel lattice every
  2x2 quad is co-circular, so the diagonal Qhull picks is arbitrary.  ``diagonals`` recovers the
  diagonals a given Delaunay run chose so that the CUDA kernel can be compared face-for-face.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

import numpy as np


def image2world(points, depth, K, R=np.eye(3), t=np.zeros((3, 1)), scale_factor=1.0):
    """hive/geometric.py:183-206."""
    num_points = points.shape[0]
    points2d = np.vstack((points.T * scale_factor, np.ones(num_points)))
    points_camera_space = np.linalg.inv(K) @ points2d
    points_world_space = R.T @ (depth * points_camera_space - t)
    return points_world_space.T


def world2image(points, K, R=np.eye(3), t=np.zeros((3, 1)), scale_factor=1.0, dtype=np.int32):
    """hive/geometric.py:155-180."""
    camera_space_coords = K @ (R @ points.T + t)
    depth = camera_space_coords[2, :]
    pixel_coords = camera_space_coords[0:2, :] / depth / scale_factor
    if issubclass(dtype, np.integer):
        pixel_coords = np.round(pixel_coords)
    pixel_coords = np.array(pixel_coords.T, dtype=dtype)
    return pixel_coords, depth


def point_cloud_from_depth(depth, mask, K, R=np.eye(3), t=np.zeros((3, 1))):
    """hive/geometric.py:107-126."""
    valid_pixels = mask & (depth > 0.0)
    V, U = valid_pixels.nonzero()
    points2d = np.array([U, V]).T
    return image2world(points2d, depth[valid_pixels], K, R, t)


def triangulate_faces(points, return_raw=False):
    """Pipeline._triangulate_faces, hive/pipeline.py:651-667."""
    from scipy.spatial import Delaunay
    tri = Delaunay(points)
    faces = np.asarray(tri.simplices)
    faces = faces[:, ::-1]  # "reverse winding order to ensure culling works as expected"
    return faces


def filter_faces(points2d, depth, faces, max_pixel_distance=2, max_depth_distance=0.1):
    """Pipeline._filter_faces, hive/pipeline.py:670-694 (options.py:274 defaults)."""
    pixel_distances = np.linalg.norm(points2d[faces[:, [0, 2, 0]]] - points2d[faces[:, [1, 1, 2]]], axis=-1)
    depth_proj = depth.reshape((*depth.shape, 1))
    depth_distances = np.linalg.norm(depth_proj[faces[:, [0, 2, 0]]] - depth_proj[faces[:, [1, 1, 2]]], axis=-1)
    valid_faces = np.all((pixel_distances <= max_pixel_distance) & (depth_distances <= max_depth_distance), axis=1)
    return faces[valid_faces]


def frame_mesh(depth, mask, K, R, t, max_pixel_distance=2, max_depth_distance=0.1):
    """process_frame's geometric core, hive/pipeline.py:386-409: vertices, filtered faces, and the raw Delaunay faces."""
    vertices = point_cloud_from_depth(depth, mask, K, R, t)
    valid = mask & (depth > 0)
    v, u = valid.nonzero()
    points2d = np.array([u, v]).T
    masked_depth = depth[valid]
    raw = triangulate_faces(points2d)
    faces = filter_faces(points2d, masked_depth, raw, max_pixel_distance, max_depth_distance)
    return vertices, faces, points2d, masked_depth, raw


def diagonals(points2d, raw_faces, H, W):
    """Per-pixel override bitmap for the CUDA kernel from the diagonals this Delaunay run chose: bit 0 = the quad
    whose top-left pixel this is was split along its anti-diagonal, bit 1 = the hole at this pixel was split N-S."""
    diag = np.zeros((H, W), dtype=np.uint8)
    p = points2d[raw_faces]  # (F, 3, 2)
    for a, b in ((0, 1), (1, 2), (2, 0)):
        d = p[:, b] - p[:, a]
        lo = np.minimum(p[:, a], p[:, b])
        third = p[:, 3 - a - b]
        unit = np.abs(p - p.mean(axis=1, keepdims=True)).max(axis=(1, 2)) < 1.0  # fits in a 2x2 quad
        anti = unit & (np.abs(d[:, 0]) == 1) & (np.abs(d[:, 1]) == 1) & (d[:, 0] * d[:, 1] == -1)
        diag[lo[anti, 1], lo[anti, 0]] |= 1
        # bridge triangles: the long edge (length 2) is the chosen diagonal of the diamond around an absent pixel;
        # only diamonds with all four neighbours present have a choice, a vertical long edge marks N-S
        vert = (d[:, 0] == 0) & (np.abs(d[:, 1]) == 2) & (np.abs(third[:, 0] - p[:, a, 0]) == 1) & \
               (third[:, 1] == lo[:, 1] + 1)
        diag[lo[vert, 1] + 1, lo[vert, 0]] |= 2
    return diag


def canonical_faces(faces):
    """Faces as a sorted array of cyclically-normalised triples (orientation preserved): set comparison."""
    f = np.asarray(faces, dtype=np.int64)
    k = np.argmin(f, axis=1)
    rolled = np.stack([f[np.arange(len(f)), (k + i) % 3] for i in range(3)], axis=1)
    return rolled[np.lexsort((rolled[:, 2], rolled[:, 1], rolled[:, 0]))]


def get_mesh_texture_and_uv(vertices, image, camera_matrix, rotation=np.eye(3), translation=np.zeros((3, 1)),
                            scale_factor=1.0):
    """hive/pipeline.py:782-810 over the oracle's world2image."""
    uv, _ = world2image(vertices, camera_matrix, rotation, translation, scale_factor)
    min_u, min_v = np.min(np.round(uv), axis=0).astype(int)
    max_u, max_v = np.max(np.round(uv), axis=0).astype(int) + 1
    texture = image[min_v:max_v, min_u:max_u, :].copy()
    uv -= np.min(np.round(uv), axis=0)
    return texture, uv


def atlas_layout(shapes, n_rows=1):
    """Placement rule of hive/pipeline.py:814-858: crops go into ``n_rows`` rows of ``len // n_rows + 1`` slots each, left
    to right; a row is as tall as its tallest crop and the canvas as wide as its widest row.  Returns
    (per-crop (x0, y0), per-row (first index, end index, y0), canvas_h, canvas_w).  A row that receives no crop makes the
    reference index past the end of the list: IndexError."""
    n = len(shapes)
    n_columns = n // n_rows + 1
    places, rows = [None] * n, []
    y0 = canvas_w = 0
    for i in range(n_rows):
        first, end = i * n_columns, min((i + 1) * n_columns, n)
        if first >= n:
            raise IndexError("list index out of range")
        x0 = row_h = 0
        for k in range(first, end):
            h, w = shapes[k][:2]
            places[k] = (x0, y0)
            x0 += w
            row_h = max(row_h, h)
        rows.append((first, end, y0))
        canvas_w = max(canvas_w, x0)
        y0 += row_h
    return places, rows, y0, canvas_w


def pack_textures(textures_atlas, uvs_atlas, n_rows=1):
    """hive/pipeline.py:810-868 without the side effect on ``uvs_atlas`` (the reference shifts the caller's arrays in
    place; both call sites, pipeline.py:469 and :619, drop them afterwards)."""
    places, rows, canvas_h, canvas_w = atlas_layout([t.shape for t in textures_atlas], n_rows)
    atlas = np.zeros((canvas_h, canvas_w, 3), dtype=np.uint8)
    shifted = []
    for k, (tex, uv) in enumerate(zip(textures_atlas, uvs_atlas)):
        uv = np.array(uv, copy=True)
        x0, y0 = places[k]
        h, w = tex.shape[:2]
        atlas[y0:y0 + h, x0:x0 + w, :] = tex
        uv[:, 0] += x0
        uv[:, 1] += y0
        shifted.append(uv)
    final_uvs = np.zeros((0, 2))
    for uv in shifted:
        final_uvs = np.vstack((final_uvs, uv))
    final_uvs[:, 0] /= canvas_w
    final_uvs[:, 1] = 1. - final_uvs[:, 1] / canvas_h
    return atlas, final_uvs


def face_adjacency(faces):
    """trimesh.Trimesh.face_adjacency as HIVE uses it: pairs of faces sharing an edge that exactly two faces use."""
    faces = np.asarray(faces)
    edges = np.sort(faces[:, [0, 1, 1, 2, 2, 0]].reshape(-1, 2), axis=1)
    face_of = np.repeat(np.arange(len(faces)), 3)
    order = np.lexsort((edges[:, 1], edges[:, 0]))
    e = edges[order]
    start = np.ones(len(e), bool)
    start[1:] = (e[1:] != e[:-1]).any(axis=1)
    run_id = np.cumsum(start) - 1
    run_len = np.bincount(run_id)
    first = np.flatnonzero(start)[run_len == 2]
    return np.stack([face_of[order[first]], face_of[order[first + 1]]], axis=1)


def cleanup_with_connected_components(vertices, faces, is_object=True, min_components=5):
    """hive/pipeline.py:741-779 with trimesh.graph.connected_components restated over scipy.sparse.csgraph (trimesh's
    own default engine): nodes = faces with a neighbour, groups ordered by label, components below min_len dropped.
    Like trimesh.Trimesh(vertices, faces) (process=True) the vertices that no input face refers to are dropped and the
    faces renumbered; the survivors keep their order here (trimesh orders them by a row hash: PARITY UNPINNED)."""
    from scipy.sparse import coo_matrix
    from scipy.sparse.csgraph import connected_components
    faces = np.asarray(faces)
    adj = face_adjacency(faces)
    mask = np.zeros(len(faces), dtype=bool)
    if len(adj):
        n = len(faces)
        graph = coo_matrix((np.ones(len(adj), bool), (adj[:, 0], adj[:, 1])), shape=(n, n))
        _, labels = connected_components(graph, directed=False)
        nodes = np.unique(adj)
        comps = {}
        for node in nodes:
            comps.setdefault(labels[node], []).append(node)
        components = [np.array(c) for _, c in sorted(comps.items()) if len(c) >= min_components]
        if components:
            if is_object:
                mask[components[int(np.argmax([len(c) for c in components]))]] = True
            else:
                mask[np.concatenate(components)] = True
    vertices = np.asarray(vertices)
    referenced = np.zeros(len(vertices), dtype=bool)
    referenced[faces.reshape(-1).astype(np.int64)] = True
    new_index = np.cumsum(referenced) - 1
    return vertices[referenced], new_index[faces[mask].astype(np.int64)].astype(faces.dtype)


def process_frame(rgb, depth, mask_encoded, pose, camera_matrix, dilation_iterations=0, max_pixel_distance=2,
                  max_depth_distance=0.1, min_num_components=5, disable_coverage_constraint=False):
    """hive/pipeline.py:340-484 for the dynamic objects (object ids >= 1), decimation and billboard off, composed from the
    restatements above.  Returns (frame_vertices, frame_faces, atlas or None, uv or None, kept ids, {id: (points2d, raw)})."""
    from . import driver_oracle as drv
    rgb = np.ascontiguousarray(rgb[:, :, :3])
    rotation, translation = pose[:3, :3], pose[:3, 3:]
    frame_vertices, frame_faces = np.zeros((0, 3)), np.zeros((0, 3))
    uv_atlas, texture_atlas, kept, raws = [], [], [], {}
    vertex_count = 0
    for object_id in range(1, int(mask_encoded.max()) + 1):
        mask = drv.dilate_mask(mask_encoded == object_id, dilation_iterations)
        if mask.mean() < 0.01 and not disable_coverage_constraint:
            continue
        vertices = point_cloud_from_depth(depth, mask, camera_matrix, rotation, translation)
        if len(vertices) < 9:
            continue
        valid = mask & (depth > 0.0)
        v, u = valid.nonzero()
        points2d = np.vstack((u, v)).T
        raw = triangulate_faces(points2d)
        raws[object_id] = (points2d, raw)
        faces = filter_faces(points2d, depth[valid], raw, max_pixel_distance, max_depth_distance)
        if len(faces) < 1:
            continue
        vertices, faces = cleanup_with_connected_components(vertices, faces, True, min_num_components)
        texture, uv = get_mesh_texture_and_uv(vertices, rgb, camera_matrix, rotation, translation)
        texture_atlas.append(texture)
        uv_atlas.append(uv)
        frame_vertices = np.vstack((frame_vertices, vertices))
        frame_faces = np.vstack((frame_faces, faces + vertex_count))
        vertex_count += len(vertices)
        kept.append(object_id)
    if not texture_atlas:
        return frame_vertices, frame_faces, None, None, kept, raws
    atlas, final_uv = pack_textures(texture_atlas, uv_atlas, n_rows=1)
    return frame_vertices, frame_faces, atlas, final_uv, kept, raws
