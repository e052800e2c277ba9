"""GPU parity of the per-frame depth-to-mesh path against the oracle (oracle/d2m_oracle.py, restating
hive/geometric.py:107-206 and hive/pipeline.py:651-694) and the golden vectors produced by the real hive.geometric.

Vertices: float64, compared to 1e-9 relative (the reference multiplies through BLAS dgemm, whose summation order is
not specified).  Faces: Qhull's diagonal on co-circular lattice quads is arbitrary, so face sets are
compared exactly after feeding the kernel the diagonals that the oracle's Delaunay run happened to choose, and through
diagonal-independent properties otherwise.
"""
import os

import numpy as np
import pytest

from hive_b200.synthetic import Scene
from oracle import d2m_oracle as d2m

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _frame(config="small", index=2):
    scene = Scene(config, "room")
    fr = scene.frame(index)
    w2c = np.linalg.inv(fr.pose)
    return scene, fr, w2c[:3, :3], w2c[:3, 3:]


def _close(a, b, rel=1e-9):
    return np.abs(a - b).max() <= rel * max(1.0, np.abs(b).max())


def test_point_cloud_matches_reference_golden():
    from hive_b200 import geometric
    g = np.load(os.path.join(GOLDEN, "backproject.npz"))
    scene, fr, R, t = _frame("small", int(g["frame_index"]))
    pts = geometric.point_cloud_from_depth(fr.depth, fr.mask > 0, scene.K, R, t)
    assert pts.dtype == np.float64 and pts.shape == g["points"].shape
    assert _close(pts, g["points"])
    full = geometric.point_cloud_from_depth(fr.depth, np.ones_like(fr.mask, dtype=bool), scene.K, R, t)
    assert len(full) == int(g["n_nomask"]) and _close(full[:64], g["points_nomask_head"])
    uv, depth = geometric.world2image(g["points"], scene.K, R, t)
    assert uv.dtype == np.int32 and np.array_equal(uv, g["uv"]) and _close(depth, g["depth"])
    # image2world on explicit points, and the float (non-rounded) branch of world2image
    v, u = (fr.mask > 0).nonzero()
    sel = fr.depth[v, u] > 0
    p2 = np.array([u[sel], v[sel]]).T
    assert _close(geometric.image2world(p2, fr.depth[v[sel], u[sel]], scene.K, R, t), g["points"])
    uvf, _ = geometric.world2image(g["points"], scene.K, R, t, dtype=np.float64)
    assert np.abs(uvf - p2).max() < 1e-4  # inv(K) is taken in float32 (geometric.py:202)


@pytest.mark.parametrize("config,index", [("small", 2), ("small", 5), ("tiny", 1)])
def test_frame_mesh_matches_oracle_with_its_diagonals(config, index):
    from hive_b200 import pipeline
    scene, fr, R, t = _frame(config, index)
    mask = fr.mask > 0
    ov, of, op2, omd, raw = d2m.frame_mesh(fr.depth, mask, scene.K, R, t)
    diag = d2m.diagonals(op2, raw, scene.H, scene.W)
    verts, faces, p2, md = pipeline.depth_to_mesh(fr.depth, mask, scene.K, R, t, diag=diag)
    assert verts.dtype == np.float64 and faces.dtype == np.int32
    assert np.array_equal(p2, op2) and np.array_equal(md, omd)
    assert _close(verts, ov)
    assert np.array_equal(d2m.canonical_faces(faces), d2m.canonical_faces(of))


def _random_case(rng, H, W, hole_p, depth_jumps):
    mask = rng.random((H, W)) > hole_p
    depth = np.full((H, W), 1.5, np.float32)
    if depth_jumps:
        depth += (rng.random((H, W)) * 0.3).astype(np.float32)  # neighbouring depths straddle the 0.1 m threshold
    depth[rng.random((H, W)) < 0.05] = 0.0
    return depth, mask


@pytest.mark.parametrize("H,W,hole_p,jumps", [(24, 31, 0.1, False), (33, 40, 0.3, True), (17, 64, 0.02, True),
                                              (40, 37, 0.5, False), (12, 12, 0.0, True)])
def test_random_masks_face_sets(H, W, hole_p, jumps):
    from hive_b200 import pipeline
    rng = np.random.default_rng(H * 1000 + W)
    depth, mask = _random_case(rng, H, W, hole_p, jumps)
    K = np.array([[50.0, 0, W / 2], [0, 50.0, H / 2], [0, 0, 1]], dtype=np.float32)
    ov, of, op2, omd, raw = d2m.frame_mesh(depth, mask, K, np.eye(3), np.zeros((3, 1)))
    diag = d2m.diagonals(op2, raw, H, W)
    verts, faces, p2, md = pipeline.depth_to_mesh(depth, mask, K, diag=diag)
    assert _close(verts, ov) and np.array_equal(p2, op2)
    assert np.array_equal(d2m.canonical_faces(faces), d2m.canonical_faces(of))
    # the production rule (no injected diagonals): every face is legal, oriented, unique
    verts2, faces2, _, _ = pipeline.depth_to_mesh(depth, mask, K)
    assert np.array_equal(verts2, verts)
    assert len(d2m.filter_faces(op2, omd, faces2)) == len(faces2)
    p = op2[faces2].astype(np.int64)
    e1, e2 = p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]
    assert (e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0] < 0).all()
    assert len(np.unique(np.sort(faces2, axis=1), axis=0)) == len(faces2)
    if not jumps:  # without depth discontinuities the face COUNT does not depend on the diagonal
        assert len(faces2) == len(of)


def test_triangulate_then_filter_composition_matches_oracle():
    from hive_b200.pipeline import MeshFilteringOptions, Pipeline
    rng = np.random.default_rng(11)
    depth, mask = _random_case(rng, 30, 45, 0.2, True)
    valid = mask & (depth > 0)
    v, u = valid.nonzero()
    p2 = np.array([u, v]).T
    md = depth[valid]
    raw = d2m.triangulate_faces(p2)
    want = d2m.filter_faces(p2, md, raw)
    diag = d2m.diagonals(p2, raw, 30, 45)
    u0, v0 = p2[:, 0].min(), p2[:, 1].min()
    sub = diag[v0:p2[:, 1].max() + 1, u0:p2[:, 0].max() + 1]
    tri = Pipeline._triangulate_faces(p2, diag=sub)
    got = Pipeline._filter_faces(p2, md, tri, MeshFilteringOptions())
    assert np.array_equal(d2m.canonical_faces(got), d2m.canonical_faces(want))
    # _filter_faces on the oracle's own (unfiltered) Delaunay faces: same faces, same order
    assert np.array_equal(Pipeline._filter_faces(p2, md, raw, MeshFilteringOptions()), want)
    # shuffled input order is mapped back to the caller's indices
    perm = rng.permutation(len(p2))
    tri_s = Pipeline._triangulate_faces(p2[perm], diag=sub)
    assert np.array_equal(d2m.canonical_faces(perm[tri_s]), d2m.canonical_faces(tri))


def test_max_pixel_distance_below_sqrt2_leaves_no_faces():
    """Every triangle of lattice points has an edge >= sqrt2, so a smaller max_pixel_distance filters all faces -- the
    reference (Delaunay + _filter_faces) agrees; the vertices are unaffected."""
    from hive_b200 import pipeline
    scene, fr, R, t = _frame("small", 1)
    w2c = np.eye(4)
    w2c[:3, :3], w2c[:3, 3:] = R, t
    mask = fr.mask > 0
    opts = pipeline.MeshFilteringOptions(max_pixel_distance=1.2)
    v, f, p2, md = pipeline.depth_to_mesh(fr.depth, mask, scene.K, w2c[:3, :3], w2c[:3, 3:], opts)
    ov, of = d2m.frame_mesh(fr.depth, mask, scene.K, w2c[:3, :3], w2c[:3, 3:], max_pixel_distance=1.2)[:2]
    assert f.shape == (0, 3) and len(of) == 0
    assert len(v) > 100 and _close(v, ov)
    v2, f2, _, _ = pipeline.depth_to_mesh(fr.depth, mask, scene.K, w2c[:3, :3], w2c[:3, 3:])
    assert np.array_equal(v, v2) and len(f2) > 0
    with pytest.raises(NotImplementedError):
        pipeline.depth_to_mesh(fr.depth, mask, scene.K, w2c[:3, :3], w2c[:3, 3:], pipeline.MeshFilteringOptions(max_pixel_distance=2.5))


def test_batch_equals_single_frames_and_handles_empty_frames():
    from hive_b200 import geometric, pipeline
    scene = Scene("small", "room")
    frames = [scene.frame(i) for i in range(4)]
    depths = np.stack([f.depth for f in frames])
    masks = np.stack([f.mask > 0 for f in frames])
    masks[2] = False  # a frame without any foreground
    poses = np.stack([np.linalg.inv(f.pose) for f in frames])
    batch = geometric.depth_to_mesh_batch(depths, masks, scene.K, poses)
    assert len(batch) == 4
    assert batch.frame_offsets[2, 0] == batch.frame_offsets[3, 0] and batch.frame_offsets[2, 1] == batch.frame_offsets[3, 1]
    for i in (0, 1, 3):
        v, f, p2, md = batch.frame(i)
        sv, sf, sp2, smd = pipeline.depth_to_mesh(depths[i], masks[i], scene.K, poses[i][:3, :3], poses[i][:3, 3:])
        assert np.array_equal(v, sv) and np.array_equal(f, sf) and np.array_equal(p2, sp2)
        assert len(v) > 0 and f.max() < len(v)
    v, f, _, _ = batch.frame(2)
    assert v.shape == (0, 3) and f.shape == (0, 3)


def test_batches_without_the_mid_call_read_equal_the_exactly_sized_call():
    """Second and later batches of a shape emit into buffers sized from the previous batch before the host has seen the
    totals (hv_d2m_emit_capped): same result when the guess is generous, and when it is far too small (repeated exactly)."""
    from hive_b200 import geometric
    scene = Scene("small", "room")
    frames = [scene.frame(i) for i in range(3)]
    depths = np.stack([f.depth for f in frames])
    poses = np.stack([np.linalg.inv(f.pose) for f in frames])
    sparse = np.stack([f.mask > 0 for f in frames])
    dense = np.ones_like(sparse)  # every pixel with a depth: many times the vertices of the instance masks

    def run(masks, fresh):
        if fresh:
            geometric._D2M_CAPS.clear()  # the exactly sized path (totals read between count and emit)
        b = geometric.depth_to_mesh_batch(depths, masks, scene.K, poses)
        return b.to_host() + (b.frame_offsets.copy(),)

    want_sparse, want_dense = run(sparse, True), run(dense, True)
    assert len(want_dense[0]) > 3 * len(want_sparse[0]) > 0
    geometric._D2M_CAPS.clear()
    for masks, want in ((sparse, want_sparse), (sparse, want_sparse), (dense, want_dense), (dense, want_dense), (sparse, want_sparse)):
        got = run(masks, False)  # 1st: exact; 2nd: guess fits; 3rd: guess far too small; 4th: fits; 5th: far too large
        for a, b in zip(got, want):
            assert a.shape == b.shape and np.array_equal(a, b)


@pytest.mark.parametrize("chunk_frames,with_masks", [(3, True), (7, True), (2, False)])
def test_pipelined_host_call_equals_the_batch(chunk_frames, with_masks):
    """depth_to_mesh_host (chunks of frames, upload / kernels / download overlapped on three streams) returns, frame by
    frame, what the one-batch call returns -- chunks that do not divide the clip, an empty frame, bool masks."""
    from hive_b200 import geometric
    scene = Scene("small", "room")
    frames = [scene.frame(i % 5) for i in range(7)]
    depths = np.stack([f.depth for f in frames])
    masks = np.stack([f.mask > 0 for f in frames])
    masks[4] = False
    poses = np.stack([np.linalg.inv(f.pose) for f in frames])
    want = geometric.depth_to_mesh_batch(depths, masks if with_masks else None, scene.K, poses)
    got = geometric.depth_to_mesh_host(depths, masks if with_masks else None, scene.K, poses, chunk_frames=chunk_frames,
                                       copy_threads=3)
    assert len(got) == 7 and got.n_vertices == want.frame_offsets[-1, 0] and got.n_faces == want.frame_offsets[-1, 1]
    for i in range(7):
        for a, b in zip(got.frame(i), want.frame(i)):
            assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b)
    if with_masks:
        assert got.frame(4)[0].shape == (0, 3)
    # a clip longer than the pinned staging budget runs part after part
    parts = geometric.depth_to_mesh_host(depths, masks if with_masks else None, scene.K, poses, chunk_frames=chunk_frames,
                                         max_pinned_bytes=1)
    assert len(parts) == 7 and parts.n_vertices == got.n_vertices
    for i in range(7):
        for a, b in zip(parts.frame(i), got.frame(i)):
            assert np.array_equal(a, b)
    # the byte-saving options: float32 vertices are the float64 ones rounded to nearest; no (u, v, depth) per vertex
    lean = geometric.depth_to_mesh_host(depths, masks if with_masks else None, scene.K, poses, chunk_frames=chunk_frames,
                                        vertex_dtype=np.float32, want_points=False)
    for i in range(7):
        v, f, p2, md = lean.frame(i)
        wv, wf, _, _ = want.frame(i)
        assert v.dtype == np.float32 and np.array_equal(v, wv.astype(np.float32)) and np.array_equal(f, wf)
        assert p2 is None and md is None


def test_texture_crop_and_uv_match_oracle():
    """Pipeline._get_mesh_texture_and_uv (hive/pipeline.py:782-810): crop + UVs bit-exact."""
    from hive_b200.pipeline import Pipeline
    from hive_b200 import geometric
    from oracle import d2m_oracle as d2m
    sc = Scene("small", "room")
    fr = sc.frame(3)
    w2c = np.linalg.inv(fr.pose)
    R, t = w2c[:3, :3], w2c[:3, 3:]
    verts = geometric.point_cloud_from_depth(fr.depth, fr.mask == 1, sc.K, R, t)
    assert len(verts) > 100
    tex, uv = Pipeline._get_mesh_texture_and_uv(verts, fr.rgb, sc.K, R, t)
    otex, ouv = d2m.get_mesh_texture_and_uv(d2m.point_cloud_from_depth(fr.depth, fr.mask == 1, sc.K, R, t), fr.rgb, sc.K, R, t)
    assert tex.shape == otex.shape and np.array_equal(tex, otex)
    assert uv.dtype == ouv.dtype and np.array_equal(uv, ouv)
    assert uv.min() == 0 and tex.shape[0] == uv[:, 1].max() + 1 and tex.shape[1] == uv[:, 0].max() + 1


def _strip(n_quads, v0):
    """A strip of n_quads quads (2 triangles each) on fresh vertices starting at index v0."""
    faces = []
    for q in range(n_quads):
        a, b, c, d = v0 + 2 * q, v0 + 2 * q + 1, v0 + 2 * q + 2, v0 + 2 * q + 3
        faces += [[a, b, c], [c, b, d]]
    return faces, v0 + 2 * n_quads + 2


@pytest.mark.parametrize("is_object", [True, False])
@pytest.mark.parametrize("min_components", [1, 5, 9])
def test_connected_components_cleanup_matches_oracle(is_object, min_components):
    """Pipeline._cleanup_with_connected_components (hive/pipeline.py:741-779) against the scipy restatement."""
    from hive_b200.pipeline import Pipeline
    faces, nv = [], 0
    for n_quads in (6, 2, 6, 1, 4):  # two largest components tie (12 faces each): the first one wins for objects
        f, nv = _strip(n_quads, nv)
        faces += f
    faces.append([nv, nv + 1, nv + 2])  # an isolated face: no neighbour, always dropped
    nv += 3
    fan, hub = [], nv  # three faces around one shared edge: that edge links nothing (trimesh requires exactly two users)
    fan += [[hub, hub + 1, hub + 2], [hub, hub + 1, hub + 3], [hub, hub + 1, hub + 4], [hub + 2, hub + 5, hub]]
    faces += fan
    nv += 6
    faces = np.array(faces, dtype=np.int32)
    rng = np.random.default_rng(3)
    faces = faces[rng.permutation(len(faces))]
    verts = rng.standard_normal((nv + 4, 3))  # the last four vertices belong to no face: they do not come back
    gv, gf = Pipeline._cleanup_with_connected_components(verts, faces, is_object=is_object, min_components=min_components)
    ov, of = d2m.cleanup_with_connected_components(verts, faces, is_object=is_object, min_components=min_components)
    assert np.array_equal(gv, verts[:nv]) and np.array_equal(gv, ov) and gv.dtype == verts.dtype
    assert np.array_equal(gf, of) and gf.dtype == faces.dtype
    if is_object and min_components <= 12:
        assert len(gf) == 12


def test_connected_components_on_a_depth_mesh():
    from hive_b200 import pipeline
    from hive_b200.pipeline import Pipeline
    sc = Scene("small", "room")
    fr = sc.frame(1)
    w2c = np.linalg.inv(fr.pose)
    R, t = w2c[:3, :3], w2c[:3, 3:]
    depth = fr.depth.copy()
    depth[::7, ::5] = 0.0  # punch holes so that fragments and pinched regions appear
    verts, faces, _, _ = pipeline.depth_to_mesh(depth, fr.mask > 0, sc.K, R, t)
    assert len(faces) > 2000
    for is_object, min_c in ((True, 5), (False, 5), (False, 40)):
        gv, gf = Pipeline._cleanup_with_connected_components(verts, faces, is_object=is_object, min_components=min_c)
        ov, of = d2m.cleanup_with_connected_components(verts, faces, is_object=is_object, min_components=min_c)
        assert np.array_equal(gf, of) and 0 < len(gf) <= len(faces)
        assert np.array_equal(gv, ov) and len(gv) == len(np.unique(faces))  # the vertices the input faces refer to
    keep, labels = pipeline.face_component_mask(faces, len(verts), 5, largest_only=False, return_labels=True)
    assert len(np.unique(labels[keep])) >= 2  # both instances survive as separate components


# ---- against the reference's own static methods (tests/golden/pipeline_statics.npz, oracle/make_pipeline_golden.py) ----

def _statics_gold():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "pipeline_statics.npz"))


def test_filter_faces_matches_reference_golden():
    """Pipeline._filter_faces (pipeline.py:670-694) on the reference's own Delaunay faces: kept faces equal, in order."""
    from hive_b200.pipeline import MeshFilteringOptions, Pipeline
    g = _statics_gold()
    got = Pipeline._filter_faces(g["ff_points2d"], g["ff_depth"], g["ff_raw"], MeshFilteringOptions())
    assert np.array_equal(got, g["ff_kept"])


def test_texture_and_uv_match_reference_golden():
    """Pipeline._get_mesh_texture_and_uv (pipeline.py:782-808): crop and UVs bit-equal to the reference's."""
    from hive_b200.pipeline import Pipeline
    from hive_b200.synthetic import Scene
    g = _statics_gold()
    sc = Scene("small", "room")
    texture, uv = Pipeline._get_mesh_texture_and_uv(g["tx_vertices"], sc.frame(2).rgb, sc.K, g["tx_R"], g["tx_t"])
    assert np.array_equal(texture, g["tx_texture"])
    assert uv.dtype == g["tx_uv"].dtype and np.array_equal(uv, g["tx_uv"])


def test_pack_textures_matches_reference_golden():
    """Pipeline._pack_textures (pipeline.py:810-868): atlas bytes and float64 UVs bit-equal on all golden layouts
    (one row, one crop, two rows, float UVs, crops of zero width / without vertices)."""
    from hive_b200.pipeline import Pipeline
    g = _statics_gold()
    for i in range(int(g["pk_cases"])):
        n = int(g[f"pk{i}_n"])
        tex = [g[f"pk{i}_tex{k}"] for k in range(n)]
        uvs = [g[f"pk{i}_uv{k}"] for k in range(n)]
        atlas, final = Pipeline._pack_textures(tex, uvs, n_rows=int(g[f"pk{i}_rows"]))
        assert atlas.dtype == np.uint8 and atlas.shape == g[f"pk{i}_atlas"].shape, i
        assert np.array_equal(atlas, g[f"pk{i}_atlas"]), i
        assert final.dtype == np.float64 and np.array_equal(final, g[f"pk{i}_final"]), i


def test_pack_textures_large_and_errors():
    """Frame-sized crops against the numpy restatement; the reference's IndexError for a row without crops."""
    from hive_b200.pipeline import Pipeline
    from oracle import d2m_oracle as d2m
    rng = np.random.default_rng(9)
    shapes = [(480, 300), (217, 640), (1, 1), (333, 129)]
    tex = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in shapes]
    uvs = [np.stack([rng.integers(0, w, 5000), rng.integers(0, h, 5000)], axis=1).astype(np.int32) for h, w in shapes]
    for n_rows in (1, 2):
        atlas, final = Pipeline._pack_textures(tex, uvs, n_rows=n_rows)
        want_atlas, want_final = d2m.pack_textures(tex, uvs, n_rows)
        assert np.array_equal(atlas, want_atlas) and np.array_equal(final, want_final)
    with pytest.raises(IndexError):
        Pipeline._pack_textures(tex, uvs, n_rows=3)  # 2 slots per row: the third row receives nothing
    with pytest.raises(IndexError):
        Pipeline._pack_textures(tex[:2], uvs[:2], n_rows=2)
    with pytest.raises(IndexError):
        Pipeline._pack_textures([], [], n_rows=1)


# ---- the composed per-frame step (pipeline.py:340-484) -----------------------------------------------------------------

def _with_small_blobs(mask, depth):
    """Adds instance 3 (a 3x3 blob: below 1 % coverage, exactly 9 vertices) and instance 4 (2x2: below 9 vertices) on
    background pixels with valid depth."""
    mask = mask.copy()
    free = (mask == 0) & (depth > 0)
    H, W = mask.shape
    placed = []
    for size, obj in ((3, 3), (2, 4)):
        for v in range(2, H - 4):
            u = next((u for u in range(2, W - 4) if free[v - 1:v + size + 1, u - 1:u + size + 1].all()), None)
            if u is not None:
                mask[v:v + size, u:u + size] = obj
                free[v - 1:v + size + 1, u - 1:u + size + 1] = False
                placed.append(obj)
                break
    assert placed == [3, 4]
    return mask


@pytest.mark.parametrize("config,index,dilation", [("small", 2, 2), ("small", 5, 0), ("tiny", 1, 1)])
@pytest.mark.parametrize("disable_coverage", [False, True])
def test_process_frame_matches_oracle(config, index, dilation, disable_coverage):
    """Objects kept / skipped, stacked vertices and faces, atlas and UVs of the whole per-frame step equal the composed
    numpy restatement (Qhull's diagonals injected per object; faces compared as sets, everything else element-wise)."""
    from hive_b200 import image_processing, pipeline
    scene, fr, R, t = _frame(config, index)
    pose = np.eye(4)
    pose[:3, :3], pose[:3, 3:] = R, t
    mask = _with_small_blobs(fr.mask, fr.depth)
    ov, of, oatlas, ouv, okept, raws = d2m.process_frame(fr.rgb, fr.depth, mask, pose, scene.K, dilation,
                                                         disable_coverage_constraint=disable_coverage)
    diags = {k: d2m.diagonals(p2, raw, scene.H, scene.W) for k, (p2, raw) in raws.items()}
    got = pipeline.process_frame(fr.rgb, fr.depth, mask, pose, scene.K,
                                 image_processing.MaskDilationOptions(num_iterations=dilation),
                                 pipeline.MeshFilteringOptions(), disable_coverage_constraint=disable_coverage, diags=diags)
    assert list(got.object_ids) == okept
    if not disable_coverage:
        assert 3 not in okept and 4 not in okept      # both blobs fall to the 1 % coverage rule
    elif dilation == 0:
        assert 3 in okept and 4 not in okept          # 3x3 = exactly 9 vertices stays, 2x2 falls to the 9-vertex rule
    assert got.vertices.dtype == np.float64 and got.faces.dtype == np.float64
    assert got.vertices.shape == ov.shape and _close(got.vertices, ov)
    assert np.array_equal(d2m.canonical_faces(got.faces), d2m.canonical_faces(of))
    assert np.array_equal(got.texture, oatlas)
    assert got.uv.dtype == np.float64 and np.array_equal(got.uv, ouv)


def test_process_frame_without_objects_is_empty():
    from hive_b200 import pipeline
    scene, fr, R, t = _frame("tiny", 1)
    pose = np.eye(4)
    pose[:3, :3], pose[:3, 3:] = R, t
    got = pipeline.process_frame(fr.rgb, fr.depth, np.zeros_like(fr.mask), pose, scene.K)
    assert got.is_empty and got.vertices.shape == (0, 3) and got.faces.shape == (0, 3)


def test_process_clip_partitions_equal_the_frame_loop():
    """process_clip over a short clip: the per-rank parts (ranks emulated one after the other on this GPU, gather=False)
    put together equal the single-process loop, which equals process_frame frame by frame."""
    from hive_b200 import image_processing, pipeline
    scene = Scene("small", "room")
    n = 5
    frames = [scene.frame(i) for i in range(n)]

    def load(i):
        w2c = np.linalg.inv(frames[i].pose)
        return frames[i].rgb, frames[i].depth, frames[i].mask, w2c

    dil = image_processing.MaskDilationOptions(num_iterations=1)
    whole = pipeline.process_clip(n, load, scene.K, dil)
    assert len(whole) == n and any(not m.is_empty for m in whole)
    for i, m in enumerate(whole):
        one = pipeline.process_frame(*load(i), scene.K, dil)
        assert np.array_equal(m.vertices, one.vertices) and np.array_equal(m.faces, one.faces) and m.object_ids == one.object_ids
    for world in (2, 3):
        got = [None] * n
        for rank in range(world):
            start, part = pipeline.process_clip(n, load, scene.K, dil, world_size=world, rank=rank, gather=False)
            assert start == pipeline.frame_range(n, world, rank)[0]
            got[start:start + len(part)] = part
        for a, b in zip(got, whole):
            assert np.array_equal(a.vertices, b.vertices) and np.array_equal(a.faces, b.faces)
            assert a.is_empty == b.is_empty and (a.is_empty or (np.array_equal(a.texture, b.texture) and np.array_equal(a.uv, b.uv)))


@pytest.mark.parametrize("config,index,dilation,disable_coverage", [("small", 1, 0, False), ("small", 2, 2, True),
                                                                    ("tiny", 3, 1, True)])
def test_process_frame_device_resident_equals_stepwise(config, index, dilation, disable_coverage):
    """The device-resident per-frame step and the composition of the reference-shaped static methods (host arrays between
    every step) give the same FrameMesh, bit for bit."""
    from hive_b200 import image_processing, pipeline
    scene, fr, R, t = _frame(config, index)
    pose = np.eye(4)
    pose[:3, :3], pose[:3, 3:] = R, t
    mask = _with_small_blobs(fr.mask, fr.depth)
    args = (fr.rgb, fr.depth, mask, pose, scene.K, image_processing.MaskDilationOptions(num_iterations=dilation),
            pipeline.MeshFilteringOptions())
    a = pipeline.process_frame(*args, disable_coverage_constraint=disable_coverage)
    b = pipeline.process_frame_stepwise(*args, disable_coverage_constraint=disable_coverage)
    assert a.object_ids == b.object_ids and a.is_empty == b.is_empty
    assert a.vertices.dtype == b.vertices.dtype and np.array_equal(a.vertices, b.vertices)
    assert a.faces.dtype == b.faces.dtype and np.array_equal(a.faces, b.faces)
    if not a.is_empty:
        assert np.array_equal(a.texture, b.texture) and np.array_equal(a.uv, b.uv)
