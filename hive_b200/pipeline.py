"""GPU drop-ins for the triangulation hot spots of ``hive/pipeline.py`` (lines 651-694) and the fused per-frame
foreground mesh step built from them (``process_frame``, pipeline.py:386-409).

``Pipeline._triangulate_faces`` / ``Pipeline._filter_faces`` keep the reference's names and signatures.  One documented
difference: the reference's ``_triangulate_faces`` returns the *full* Delaunay triangulation (including the long
triangles that fill the convex hull) and relies on ``_filter_faces`` to drop every face with an edge longer than
``max_pixel_distance``; here ``_triangulate_faces`` already returns only the faces that can survive that pixel filter
(on the pixel lattice they have a closed form), so ``_filter_faces(_triangulate_faces(p), ...)`` -- the
only way HIVE uses them (pipeline.py:404-409, 580-582) -- is unchanged, modulo the diagonal of co-circular quads, which
Qhull picks arbitrarily.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np

from . import _lib
from . import _xfer
from . import geometric

__all__ = ["MeshFilteringOptions", "Pipeline", "depth_to_mesh", "FrameMesh", "process_frame", "process_frame_stepwise",
           "process_clip", "frame_range"]


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


@dataclass
class MeshFilteringOptions:
    """hive/options.py:274 defaults."""
    max_pixel_distance: int = 2
    max_depth_distance: float = 0.1
    min_num_components: int = 5


def depth_to_mesh(depth, mask, K, R=geometric._EYE, t=geometric._ZERO_T, options: MeshFilteringOptions = None,
                  diag=None):
    """One frame of pipeline.py:386-409 fused: returns (vertices (N,3) f64, faces (F,3) i32, points2d (N,2) i32,
    masked_depth (N,) f32)."""
    options = options or MeshFilteringOptions()
    pose = np.eye(4)
    pose[:3, :3] = R
    pose[:3, 3:] = np.asarray(t).reshape(3, 1)
    batch = geometric.depth_to_mesh_batch(np.asarray(depth)[None], None if mask is None else np.asarray(mask)[None], K,
                                          pose[None], options.max_pixel_distance, options.max_depth_distance,
                                          None if diag is None else np.asarray(diag)[None])
    return batch.frame(0)


class Pipeline:
    """Only the static hot-path helpers of ``hive.pipeline.Pipeline``."""

    @staticmethod
    def _triangulate_faces(points: np.ndarray, max_pixel_distance=2, diag=None) -> np.ndarray:
        """pipeline.py:651-667 for points on the integer pixel lattice (what HIVE passes, pipeline.py:393-397)."""
        points = np.asarray(points)
        geometric.validate_shape(points, 'points', (None, 2))
        assert np.issubdtype(points.dtype, np.integer) or np.array_equal(points, np.round(points)), \
            "hive_b200 triangulates pixel-lattice points only"
        torch = _lib.require_cuda()
        pts = points.astype(np.int64)
        u0, v0 = pts[:, 0].min(), pts[:, 1].min()
        W, H = int(pts[:, 0].max() - u0 + 1), int(pts[:, 1].max() - v0 + 1)
        lin = (pts[:, 1] - v0) * W + (pts[:, 0] - u0)
        assert len(np.unique(lin)) == len(lin), "duplicate points"
        grid = np.zeros(H * W, dtype=np.float32)
        grid[lin] = 1.0
        dev = torch.device("cuda", torch.cuda.current_device())
        d = torch.from_numpy(grid.reshape(1, H, W)).to(dev)
        g = None
        if diag is not None:
            g = torch.from_numpy(np.ascontiguousarray(diag, dtype=np.uint8).reshape(1, H, W)).to(dev)
        cams = torch.zeros((1, 12), dtype=torch.float64, device=dev)
        out = geometric.depth_to_mesh_device(d, None, np.eye(3), cams, np.float32(np.inf),
                                             geometric.d2m_flags(max_pixel_distance), g, want_points=False)
        faces = _xfer.to_host(out.faces)
        order = np.argsort(lin, kind="stable")  # rank in row-major order -> index in the caller's array
        if not np.array_equal(order, np.arange(len(lin))):
            faces = order[faces].astype(np.int32)
        return faces

    @staticmethod
    def _filter_faces(points2d, depth, faces, options: MeshFilteringOptions):
        """pipeline.py:670-694 for arbitrary faces (order preserving)."""
        points2d, depth, faces = np.asarray(points2d), np.asarray(depth), np.asarray(faces)
        geometric.validate_shape(points2d, 'points2d', (None, 2))
        geometric.validate_shape(depth, 'depth', (points2d.shape[0],))
        geometric.validate_shape(faces, 'faces', (None, 3))
        n_faces = faces.shape[0]
        if n_faces == 0:
            return faces
        torch = _lib.require_cuda()
        lib = _lib.load()
        dev = torch.device("cuda", torch.cuda.current_device())
        is64 = depth.dtype == np.float64
        p = _xfer.to_device(points2d, dev, np.int32)
        d = _xfer.to_device(depth, dev, np.float64 if is64 else np.float32)
        f = _xfer.to_device(faces, dev, np.int32)
        nbytes = ctypes.c_int64(0)
        _lib.check(lib.hv_filter_faces_scratch_bytes(n_faces, ctypes.byref(nbytes)))
        scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=dev)
        sptr = ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256)
        totals = torch.zeros(2, dtype=torch.int64, device=dev)
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        mdd = float(np.float32(options.max_depth_distance)) if not is64 else float(options.max_depth_distance)
        args = (_ptr(p), _ptr(d), int(is64), _ptr(f), n_faces, float(options.max_pixel_distance), mdd, sptr,
                int(nbytes.value))
        _lib.check(lib.hv_filter_faces(*args, 0, _ptr(totals), None, stream))
        kept = int(totals[0].item())
        out = torch.empty((kept, 3), dtype=torch.int32, device=dev)
        if kept:
            _lib.check(lib.hv_filter_faces(*args, 1, _ptr(totals), _ptr(out), stream))
        return _xfer.to_host(out).astype(faces.dtype, copy=False)

    @staticmethod
    def _get_mesh_texture_and_uv(vertices, image, camera_matrix, rotation=geometric._EYE, translation=geometric._ZERO_T,
                                 scale_factor=1.0):
        """pipeline.py:782-810: project the vertices (GPU ``world2image``, int32 pixels like the reference's default),
        crop the image to their bounding box and shift the UVs to the crop.  Returns (texture (h, w, 3), uv (N, 2))."""
        vertices, image = np.asarray(vertices), np.asarray(image)
        geometric.validate_shape(vertices, 'vertices', (None, 3))
        geometric.validate_shape(image, 'image', (None, None, 3))
        geometric.validate_camera_parameter_shapes(camera_matrix, rotation, translation)
        uv, _ = geometric.world2image(vertices, camera_matrix, rotation, translation, scale_factor)
        lo = np.min(np.round(uv), axis=0)
        min_u, min_v = lo.astype(int)
        max_u, max_v = np.max(np.round(uv), axis=0).astype(int) + 1
        texture = image[min_v:max_v, min_u:max_u, :].copy()
        uv -= lo
        return texture, uv

    @staticmethod
    def _pack_textures(textures_atlas, uvs_atlas, n_rows=1):
        """pipeline.py:810-868: the crops of a frame's objects side by side on one black canvas (``n_rows`` rows of
        ``len // n_rows + 1`` slots, row height = tallest crop, canvas width = widest row) and their UVs shifted to the
        crop's place and normalised (``u / canvas_w``, ``1 - v / canvas_h``).  Returns (atlas (H, W, 3) uint8,
        final_uvs (sum N_k, 2) float64).

        The placement is a few integers per crop (host); pixels and UVs move on the GPU (``hv_atlas_pack``).  One
        documented difference: the reference also shifts the caller's ``uvs_atlas`` arrays in place as a side effect
        (pipeline.py:837, 857); both of its call sites drop them right after, and here they are left untouched."""
        torch = _lib.require_cuda()
        lib = _lib.load()
        n = len(textures_atlas)
        assert len(uvs_atlas) == n, "one uv array per texture"
        n_columns = n // n_rows + 1
        textures = [np.asarray(t) for t in textures_atlas]
        for t in textures:
            geometric.validate_shape(t, 'texture', (None, None, 3))
        desc = np.zeros((max(n, 1), 9), dtype=np.int64)  # src byte offset, x0, y0, w, h, first uv row, src pitch, uv shift
        row_first, row_y0 = np.zeros(n_rows + 1, dtype=np.int64), np.zeros(n_rows + 1, dtype=np.int64)
        src_off = uv_row = canvas_h = canvas_w = 0
        for i in range(n_rows):
            first, end = i * n_columns, min((i + 1) * n_columns, n)
            if first >= n:  # the reference indexes textures_atlas[i * n_columns] of a row that received no crop
                raise IndexError("list index out of range")
            x0 = row_h = 0
            for k in range(first, end):
                h, w = textures[k].shape[:2]
                desc[k] = (src_off, x0, canvas_h, w, h, uv_row, w, 0, 0)
                src_off += h * w * 3
                uv_row += len(uvs_atlas[k])
                x0 += w
                row_h = max(row_h, h)
            row_first[i], row_y0[i] = first, canvas_h
            canvas_w = max(canvas_w, x0)
            canvas_h += row_h
        row_first[n_rows], row_y0[n_rows] = n, canvas_h

        dev = torch.device("cuda", torch.cuda.current_device())
        src_h = np.concatenate([t.reshape(-1) for t in textures]) if src_off else np.zeros(0, dtype=np.uint8)
        uv_h = np.vstack([np.zeros((0, 2))] + [np.asarray(u) for u in uvs_atlas])
        with torch.cuda.device(dev):
            src = _xfer.to_device(src_h, dev, np.uint8)
            uv = _xfer.to_device(uv_h, dev, np.float64)
            d_desc = _xfer.to_device(desc, dev, np.int64)
            d_rows = _xfer.to_device(np.concatenate([row_first, row_y0]), dev, np.int64)
            atlas = torch.empty((canvas_h, canvas_w, 3), dtype=torch.uint8, device=dev)
            uv_out = torch.empty((uv_row, 2), dtype=torch.float64, device=dev)
            _lib.check(lib.hv_atlas_pack(_ptr(src) if src_off else None, _ptr(d_desc), n, _ptr(d_rows),
                                         ctypes.c_void_p(d_rows.data_ptr() + 8 * (n_rows + 1)), n_rows, canvas_h, canvas_w,
                                         _ptr(atlas) if atlas.numel() else None, _ptr(uv) if uv_row else None, uv_row,
                                         _ptr(uv_out) if uv_row else None,
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
            return _xfer.to_host(atlas, uv_out)

    @staticmethod
    def _cleanup_with_connected_components(vertices, faces, is_object=True, min_components=5):
        """pipeline.py:741-779: drop floating fragments.  Faces adjacent across an edge that exactly two faces share form
        a graph; components with fewer than ``min_components`` faces go; for an object only the largest survives.

        Returns (vertices, kept faces in their original order).  Like the reference's ``trimesh.Trimesh(vertices, faces)``
        (built with ``process=True``) the returned vertex array holds only the vertices that the *input* faces refer to
        -- a valid pixel whose faces were all filtered out does not reach the texture crop -- and the faces are
        renumbered accordingly.  PARITY UNPINNED (trimesh is on neither box): trimesh additionally merges duplicate
        vertices (there are none in a depth-to-mesh result) and orders the surviving vertices by a hash of their rows;
        here they keep their original order.  A tie between equally large components goes to the one containing the
        lowest face index (trimesh: the first in its own traversal order)."""
        vertices, faces = np.asarray(vertices), np.asarray(faces)
        geometric.validate_shape(vertices, 'vertices', (None, 3))
        geometric.validate_shape(faces, 'faces', (None, 3))
        if faces.shape[0] == 0:
            return vertices, faces
        torch = _lib.require_cuda()
        lib = _lib.load()
        dev = torch.device("cuda", torch.cuda.current_device())
        n_verts, n_faces = int(vertices.shape[0]), int(faces.shape[0])
        keep = face_component_mask(faces, n_verts, min_components, largest_only=is_object)
        kept_faces = np.ascontiguousarray(faces[keep])
        with torch.cuda.device(dev):
            f_all = _xfer.to_device(faces, dev, np.int32)
            f_keep = _xfer.to_device(kept_faces, dev, np.int32) if len(kept_faces) else None
            v_in = _xfer.to_device(vertices, dev, np.float64)
            v_out = torch.empty((n_verts, 3), dtype=torch.float64, device=dev)
            f_out = torch.empty((len(kept_faces), 3), dtype=torch.int32, device=dev)
            count = torch.zeros(2, dtype=torch.int64, device=dev)
            nbytes = ctypes.c_int64(0)
            _lib.check(lib.hv_cc_compact_scratch_bytes(n_verts, ctypes.byref(nbytes)))
            scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=dev)
            sptr = ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256)
            _lib.check(lib.hv_cc_compact_vertices(_ptr(f_all), n_faces, _ptr(f_keep), len(kept_faces), _ptr(v_in), n_verts,
                                                  _ptr(v_out), _ptr(f_out) if len(kept_faces) else None, _ptr(count), sptr,
                                                  int(nbytes.value), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
            n_kept = int(count[0].item())
            out_v, out_f = _xfer.to_host(v_out[:n_kept], f_out)
        return out_v.astype(vertices.dtype, copy=False), out_f.astype(faces.dtype, copy=False)


def face_component_mask(faces, n_verts, min_len=5, largest_only=True, return_labels=False):
    """Boolean keep-mask over ``faces`` (host (F, 3) int array) from the GPU connected-components kernels."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    n_faces = int(faces.shape[0])
    f = _xfer.to_device(faces, dev, np.int32)
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        nbytes = ctypes.c_int64(0)
        _lib.check(lib.hv_cc_scratch_bytes(n_faces, ctypes.byref(nbytes)))
        scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=dev)
        sptr = ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256)
        keep = torch.empty(n_faces, dtype=torch.uint8, device=dev)
        labels = torch.empty(n_faces, dtype=torch.int32, device=dev) if return_labels else None
        _lib.check(lib.hv_cc_mask(_ptr(f), n_faces, int(n_verts), int(min_len), int(bool(largest_only)), sptr,
                                  int(nbytes.value), _ptr(keep), _ptr(labels), stream))
    if return_labels:
        k, lab = _xfer.to_host(keep, labels)
        return k.view(np.bool_), lab.view(np.uint32)
    return _xfer.to_host(keep).view(np.bool_)


@dataclass
class FrameMesh:
    """What ``process_frame`` hands to ``trimesh.Trimesh(frame_vertices, frame_faces, TextureVisuals(uv, texture))``
    (pipeline.py:472-482).  ``faces`` is float64 like the reference's (it stacks onto ``np.zeros((0, 3))``)."""
    vertices: np.ndarray
    faces: np.ndarray
    texture: np.ndarray = None
    uv: np.ndarray = None
    object_ids: tuple = ()

    @property
    def is_empty(self):
        return self.texture is None


def process_frame_stepwise(rgb, depth, mask_encoded, pose, camera_matrix, dilation_options=None, filtering_options=None,
                           disable_coverage_constraint=False, diags=None) -> FrameMesh:
    """``process_frame`` composed from the reference-shaped static methods, every one of which takes and returns host arrays
    (five host <-> device round trips per object).  Kept as the A/B partner of ``process_frame`` (same result, tests).

    The per-frame foreground mesh step, ``process_frame`` of pipeline.py:340-484, for the dynamic objects
    (``include_background=False``): for every instance id in the mask -- dilate its mask, skip it below 1 % coverage,
    back-project + triangulate + filter (one fused CUDA call), skip below 9 vertices or without faces, drop floaters
    (largest face component), crop the texture and the UVs -- then pack the crops into the frame's atlas.

    Every step is the GPU drop-in of this package; the loop, the three skip rules and the stacking are the reference's.
    Not performed: mesh decimation (``_decimate_mesh`` is OpenMesh's sequential edge collapse, out of scope: the meshes
    are passed on undecimated) and the optional billboard flattening.  ``diags``: per-object diagonal bitmaps for the
    co-circular lattice quads (tests inject the oracle's Qhull choices; ``None`` = this package's rule).
    """
    from . import image_processing
    dilation_options = dilation_options or image_processing.MaskDilationOptions(num_iterations=0)
    filtering_options = filtering_options or MeshFilteringOptions()
    rgb = np.ascontiguousarray(np.asarray(rgb)[:, :, :3])
    depth, mask_encoded = np.asarray(depth), np.asarray(mask_encoded)
    rotation, translation = geometric.get_pose_components(np.asarray(pose))

    frame_vertices = np.zeros((0, 3))
    frame_faces = np.zeros((0, 3))
    uv_atlas, texture_atlas, kept_ids = [], [], []
    vertex_count = 0
    for object_id in range(1, int(mask_encoded.max()) + 1):
        mask = image_processing.dilate_mask(mask_encoded == object_id, dilation_options)
        coverage_ratio = np.count_nonzero(mask) / mask.size  # == mask.mean()
        if coverage_ratio < 0.01 and not disable_coverage_constraint:
            continue
        vertices, faces, _, _ = depth_to_mesh(depth, mask, camera_matrix, rotation, translation, filtering_options,
                                              None if diags is None else diags.get(object_id))
        if len(vertices) < 9 or len(faces) < 1:
            continue
        vertices, faces = Pipeline._cleanup_with_connected_components(vertices, faces, True,
                                                                     min_components=filtering_options.min_num_components)
        texture, uv = Pipeline._get_mesh_texture_and_uv(vertices, rgb, camera_matrix, rotation, translation)
        texture_atlas.append(texture)
        uv_atlas.append(uv)
        frame_vertices = np.vstack((frame_vertices, vertices))
        frame_faces = np.vstack((frame_faces, faces + vertex_count))
        vertex_count += len(vertices)
        kept_ids.append(object_id)
    if len(texture_atlas) == 0:
        return FrameMesh(frame_vertices, frame_faces)
    packed_textures, packed_uv = Pipeline._pack_textures(texture_atlas, uv_atlas, n_rows=1)
    return FrameMesh(frame_vertices, frame_faces, packed_textures, packed_uv, tuple(kept_ids))


def frame_range(n_frames: int, world_size: int, rank: int):
    """Contiguous, balanced frame range ``[start, stop)`` of ``rank``: the first ``n_frames % world_size`` ranks take one
    frame more.  (Contiguous so that a rank reads one stretch of the dataset, like one worker of the reference's pool.)"""
    assert world_size >= 1 and 0 <= rank < world_size and n_frames >= 0
    base, extra = divmod(int(n_frames), int(world_size))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def process_clip(n_frames, load_frame, camera_matrix, dilation_options=None, filtering_options=None,
                 disable_coverage_constraint=False, world_size=1, rank=0, dist=None, gather=True, frame_fn=None):
    """The foreground-mesh loop of a clip, ``Pipeline._create_scene`` of pipeline.py:486-497 (``tqdm_imap(process_frame,
    range(num_frames))``: a ThreadPool over frames, hive/utils.py:93) partitioned over GPUs by frame: rank r processes
    the contiguous range ``frame_range(n_frames, world_size, r)`` on its own GPU -- no collective while it computes, the
    frames are independent -- and the variable-length results are either left where they are (``gather=False``: returns
    ``(start, [FrameMesh, ...])`` of this rank) or gathered to rank 0 by ONE collective (``gather=True``: rank 0 returns the
    ``n_frames`` meshes in frame order, the other ranks ``None``).

    ``load_frame(i) -> (rgb, depth, mask_encoded, pose)`` reads frame i of the dataset (``dataset.rgb_dataset[i]`` ... of
    pipeline.py:341-348); it is only called for this rank's frames.  ``frame_fn``: the per-frame step, ``process_frame``
    of this module unless given (the CPU tests of the partition / gather logic pass a stand-in).
    """
    frame_fn = frame_fn or process_frame
    start, stop = frame_range(n_frames, world_size, rank)
    meshes = []
    for i in range(start, stop):
        rgb, depth, mask_encoded, pose = load_frame(i)
        meshes.append(frame_fn(rgb, depth, mask_encoded, pose, camera_matrix, dilation_options, filtering_options,
                               disable_coverage_constraint))
    if world_size == 1:
        return meshes if gather else (start, meshes)
    if not gather:
        return start, meshes
    assert dist is not None, "process_clip: pass torch.distributed as `dist` when world_size > 1"
    parts = [None] * world_size if rank == 0 else None
    dist.gather_object((start, meshes), parts, dst=0)  # one variable-length gather
    if rank != 0:
        return None
    out = [None] * n_frames
    for part_start, part in parts:
        out[part_start:part_start + len(part)] = part
    assert all(m is not None for m in out), "process_clip: a frame is missing from the gather"
    return out


def process_frame(rgb, depth, mask_encoded, pose, camera_matrix, dilation_options=None, filtering_options=None,
                  disable_coverage_constraint=False, diags=None) -> FrameMesh:
    """The per-frame foreground mesh step, ``process_frame`` of pipeline.py:340-484, for the dynamic objects
    (``include_background=False``), with the objects' data resident in HBM from the mask to the atlas: all object masks are
    dilated in one batch, their coverage counted on the device, the surviving objects meshed in one batched depth-to-mesh
    call, cleaned (components + vertex compaction) and projected on the device, and the atlas is cut straight out of the
    frame.  The host sees four small read-backs per FRAME (coverage counts, mesh sizes, cleaned sizes, crop boxes: a few
    integers per object -- every one decides what is launched next) and one download of the result; the step-wise
    composition (``process_frame_stepwise``) moves every object's arrays across PCIe five times.  Same skip rules, same
    result.  Not performed: mesh decimation (OpenMesh's sequential edge collapse) and the billboard flattening.
    """
    from . import image_processing
    torch = _lib.require_cuda()
    lib = _lib.load()
    dilation_options = dilation_options or image_processing.MaskDilationOptions(num_iterations=0)
    filtering_options = filtering_options or MeshFilteringOptions()
    rgb = np.ascontiguousarray(np.asarray(rgb)[:, :, :3])
    depth, mask_encoded = np.asarray(depth), np.asarray(mask_encoded)
    H, W = depth.shape
    rotation, translation = geometric.get_pose_components(np.asarray(pose))
    empty = FrameMesh(np.zeros((0, 3)), np.zeros((0, 3)))
    n_obj = int(mask_encoded.max())
    if n_obj < 1:
        return empty
    n_iter = image_processing._check_filter(dilation_options)
    dev = torch.device("cuda", torch.cuda.current_device())
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        depth_d = _xfer.to_device(depth[None], dev, np.float32)
        rgb_d = _xfer.to_device(rgb, dev, np.uint8)
        # ---- masks of all objects: dilate (one batch), coverage (pipeline.py:370-381)
        ids = np.arange(1, n_obj + 1)
        masks_d = image_processing.dilate_masks_device(
            _xfer.to_device(mask_encoded[None] == ids[:, None, None], dev, np.uint8), n_iter)
        counts = torch.empty(n_obj, dtype=torch.int64, device=dev)
        _lib.check(lib.hv_count_nonzero(_ptr(masks_d), n_obj, H * W, _ptr(counts), stream))
        coverage = counts.cpu().numpy() / float(H * W)                                             # read-back 1
        cand = [k for k in range(n_obj) if disable_coverage_constraint or not coverage[k] < 0.01]
        if not cand:
            return empty
        # ---- back-projection + triangulation + face filter of the candidates, one batched call (pipeline.py:386-409)
        B = len(cand)
        sel = masks_d if B == n_obj else masks_d[torch.as_tensor(cand, device=dev)]
        diag_d = None
        if diags is not None:
            diag_h = np.stack([np.asarray(diags[int(ids[k])], dtype=np.uint8).reshape(H, W) if int(ids[k]) in diags
                               else np.zeros((H, W), np.uint8) for k in cand])
            diag_d = _xfer.to_device(diag_h, dev, np.uint8)
        cam = np.concatenate([np.asarray(rotation, dtype=np.float64).reshape(9), np.asarray(translation, dtype=np.float64).reshape(3)])
        cams_d = _xfer.to_device(np.tile(cam, (B, 1)), dev, np.float64)
        batch = geometric.depth_to_mesh_device(depth_d.expand(B, H, W).contiguous(), sel, geometric._kinv64(camera_matrix),
                                               cams_d, filtering_options.max_depth_distance,
                                               geometric.d2m_flags(filtering_options.max_pixel_distance), diag_d,
                                               want_points=False)                                  # read-back 2
        off = batch.frame_offsets
        alive = [j for j in range(B) if off[j + 1, 0] - off[j, 0] >= 9 and off[j + 1, 1] - off[j, 1] >= 1]
        if not alive:
            return empty
        # ---- floater removal + vertex compaction per object, on its slice of the batch (pipeline.py:741-779)
        parts, counts4 = [], torch.zeros((len(alive), 4), dtype=torch.int64, device=dev)
        for i, j in enumerate(alive):
            v0, f0 = (int(v) for v in off[j])
            nv, nf = int(off[j + 1, 0]) - v0, int(off[j + 1, 1]) - f0
            nbytes = ctypes.c_int64(0)
            _lib.check(lib.hv_cc_cleanup_scratch_bytes(nf, nv, ctypes.byref(nbytes)))
            scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=dev)
            sptr = ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256)
            v_out = torch.empty((nv, 3), dtype=torch.float64, device=dev)
            f_out = torch.empty((nf, 3), dtype=torch.int32, device=dev)
            _lib.check(lib.hv_cc_cleanup(_ptr(batch.faces[f0:f0 + nf]), nf, _ptr(batch.verts[v0:v0 + nv]), nv,
                                         int(filtering_options.min_num_components), 1, _ptr(v_out), _ptr(f_out),
                                         _ptr(counts4[i]), sptr, int(nbytes.value), stream))
            parts.append((v_out, f_out, scratch))
        sizes = counts4.cpu().numpy()                                                              # read-back 3
        # ---- projection into the frame, crop box (pipeline.py:782-808)
        cam21 = torch.from_numpy(geometric._cam21(np.asarray(camera_matrix, dtype=np.float64), rotation, translation)).to(dev)
        boxes = torch.empty((len(alive), 4), dtype=torch.int64, device=dev)
        uvs = []
        for i, (v_out, f_out, _) in enumerate(parts):
            nk = int(sizes[i, 0])
            uv = torch.empty((nk, 2), dtype=torch.float64, device=dev)
            zc = torch.empty((nk,), dtype=torch.float64, device=dev)
            _lib.check(lib.hv_world2image(_ptr(v_out), nk, _ptr(cam21), 1.0, 1, _ptr(uv), _ptr(zc), stream))
            _lib.check(lib.hv_uv_bbox(_ptr(uv), nk, _ptr(boxes[i]), stream))
            uvs.append(uv)
        box = boxes.cpu().numpy()                                                                  # read-back 4
        # ---- atlas (pipeline.py:810-868 with n_rows = 1): layout on the host, pixels straight out of the frame
        n = len(alive)
        desc = np.zeros((n, 9), dtype=np.int64)
        x0 = canvas_h = uv_row = 0
        for i in range(n):
            min_u, min_v, max_u, max_v = (int(v) for v in box[i])
            assert min_u >= 0 and min_v >= 0, "a vertex projects left of / above the frame"
            w, h = max(min(max_u + 1, W) - min_u, 0), max(min(max_v + 1, H) - min_v, 0)  # numpy slicing clips at the frame
            desc[i] = ((min_v * W + min_u) * 3, x0, 0, w, h, uv_row, W, min_u, min_v)
            x0 += w
            canvas_h = max(canvas_h, h)
            uv_row += int(sizes[i, 0])
        canvas_w = x0
        rows = np.array([0, n, 0, canvas_h], dtype=np.int64)  # row_first, row_y0 of the single row
        d_desc, d_rows = _xfer.to_device(desc, dev, np.int64), _xfer.to_device(rows, dev, np.int64)
        uv_all = torch.cat(uvs)
        atlas = torch.empty((canvas_h, canvas_w, 3), dtype=torch.uint8, device=dev)
        uv_out = torch.empty((uv_row, 2), dtype=torch.float64, device=dev)
        _lib.check(lib.hv_atlas_pack(_ptr(rgb_d), _ptr(d_desc), n, _ptr(d_rows), ctypes.c_void_p(d_rows.data_ptr() + 16), 1,
                                     canvas_h, canvas_w, _ptr(atlas) if atlas.numel() else None, _ptr(uv_all), uv_row,
                                     _ptr(uv_out), stream))
        verts_all = torch.cat([parts[i][0][:int(sizes[i, 0])] for i in range(n)])
        faces_dev = [parts[i][1][:int(sizes[i, 2])] for i in range(n)]
        host = _xfer.to_host(verts_all, atlas, uv_out, *faces_dev)                                 # the download
    frame_vertices, packed_textures, packed_uv = host[0], host[1], host[2]
    frame_faces, vertex_count = np.zeros((0, 3)), 0
    for i in range(n):
        frame_faces = np.vstack((frame_faces, host[3 + i] + vertex_count))  # float64 like the reference's stacking
        vertex_count += int(sizes[i, 0])
    kept_ids = tuple(int(ids[cand[j]]) for j in alive)
    return FrameMesh(frame_vertices, frame_faces, packed_textures, packed_uv, kept_ids)
