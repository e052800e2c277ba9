"""Host-side pieces that need no GPU: the layout of the shared whole-mesh buffer and the per-frame accessor of the chunked
host-to-host depth-to-mesh result."""
import numpy as np
import pytest

from hive_b200 import geometric, peer


@pytest.mark.parametrize("nv,nt", [(0, 0), (1, 1), (633101, 1266202), (840538, 1680716)])
def test_mesh_layout_arrays_are_aligned_and_disjoint(nv, nt):
    offs, total = peer.mesh_layout(nv, nt)
    sizes = {"verts": 12 * nv, "normals": 12 * nv, "colors": 3 * nv, "faces": 12 * nt}
    spans = sorted((offs[k], offs[k] + sizes[k]) for k in sizes)
    assert all(o % 256 == 0 for o, _ in spans) and total % 256 == 0 and total >= max(256, spans[-1][1])
    for (a0, a1), (b0, b1) in zip(spans[:-1], spans[1:]):
        assert a1 <= b0, "arrays overlap"
    # every rank derives the same layout from the same totals
    assert peer.mesh_layout(nv, nt) == (offs, total)


def test_host_depth_meshes_frame_accessor_maps_frames_to_chunks():
    rng = np.random.default_rng(3)
    per_frame = [(5, 7), (0, 0), (3, 2), (9, 11), (1, 0)]  # (vertices, faces) of five frames in chunks of 2, 2, 1
    chunks, frame = [], 0
    for size in (2, 2, 1):
        counts = per_frame[frame:frame + size]
        off = np.zeros((size + 1, 2), dtype=np.int64)
        off[1:] = np.cumsum(np.array(counts, dtype=np.int64).reshape(size, 2), axis=0)
        nv, nf = int(off[-1, 0]), int(off[-1, 1])
        chunks.append((frame, rng.random((nv, 3)), rng.integers(0, 9, (nf, 3)).astype(np.int32),
                       rng.integers(0, 640, (nv, 2)).astype(np.int32), rng.random(nv).astype(np.float32), off))
        frame += size
    meshes = geometric.HostDepthMeshes(chunks, 5)
    assert len(meshes) == 5 and meshes.n_vertices == sum(v for v, _ in per_frame) and meshes.n_faces == sum(f for _, f in per_frame)
    for i, (nv, nf) in enumerate(per_frame):
        v, f, p2, d = meshes.frame(i)
        assert v.shape == (nv, 3) and f.shape == (nf, 3) and p2.shape == (nv, 2) and d.shape == (nv,)
    first, verts = chunks[1][0], chunks[1][1]
    assert np.array_equal(meshes.frame(first + 1)[0], verts[3:12])  # second frame of the second chunk: its rows 3..11
    with pytest.raises(IndexError):
        meshes.frame(5)
    lean = geometric.HostDepthMeshes([c[:3] + (None, None) + c[5:] for c in chunks], 5)  # want_points=False
    assert lean.frame(3)[2] is None and lean.frame(3)[3] is None and lean.frame(3)[0].shape == (9, 3)


def test_d2m_flags_ranges():
    assert geometric.d2m_flags(0.5) == 2 and geometric.d2m_flags(1.4) == 2   # below sqrt2: vertices only
    assert geometric.d2m_flags(np.sqrt(2.0)) == 0 and geometric.d2m_flags(1.9) == 0
    assert geometric.d2m_flags(2) == 1 and geometric.d2m_flags(2.2) == 1
    with pytest.raises(NotImplementedError):
        geometric.d2m_flags(np.sqrt(5.0))
