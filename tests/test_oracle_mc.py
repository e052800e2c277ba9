"""CPU checks of the marching-cubes restatement (oracle/mc_oracle.c; skimage's Lewiner code is not on the box, so
parity with it is UNPINNED -- these are the cross-checks SURVEY.md §8(c) lists instead):

* every cube configuration without an ambiguous face against the classic tables that ship in the reference tree
  (third_party/colmap/lib/PoissonRecon/MarchingCubes.cpp -> tests/golden/mc_classic_cases.npz, made by
  oracle/make_mc_classic_golden.py): same triangle count, same split into connected patches;
* closed oriented 2-manifold on noise volumes (every ambiguous case occurs), Euler characteristic of a sphere and a
  torus, every vertex on a sign-changing lattice edge, normals pointing down-hill.
"""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FACES = [[(0, y, z) for y, z in ((0, 0), (1, 0), (1, 1), (0, 1))], [(1, y, z) for y, z in ((0, 0), (1, 0), (1, 1), (0, 1))],
         [(x, 0, z) for x, z in ((0, 0), (1, 0), (1, 1), (0, 1))], [(x, 1, z) for x, z in ((0, 0), (1, 0), (1, 1), (0, 1))],
         [(x, y, 0) for x, y in ((0, 0), (1, 0), (1, 1), (0, 1))], [(x, y, 1) for x, y in ((0, 0), (1, 0), (1, 1), (0, 1))]]


def _cube_volume(cfg):
    vol = np.ones((2, 2, 2), np.float32)
    for i in range(8):
        if (cfg >> i) & 1:
            vol[i & 1, (i >> 1) & 1, (i >> 2) & 1] = -1.0
    return vol


def _has_ambiguous_face(vol):
    for f in FACES:
        s = [vol[p] > 0 for p in f]
        if s[0] == s[2] and s[1] == s[3] and s[0] != s[1]:
            return True
    return False


def _patch_sizes(faces):
    parent = list(range(len(faces)))

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a
    for i in range(len(faces)):
        for j in range(i):
            if set(faces[i]) & set(faces[j]):
                parent[find(i)] = find(j)
    sizes = {}
    for i in range(len(faces)):
        sizes[find(i)] = sizes.get(find(i), 0) + 1
    return sorted(sizes.values(), reverse=True)


def test_unambiguous_configurations_match_the_classic_tables(c_oracle):
    g = np.load(os.path.join(GOLDEN, "mc_classic_cases.npz"))
    checked = 0
    for cfg in range(1, 255):
        vol = _cube_volume(cfg)
        if _has_ambiguous_face(vol):
            continue
        verts, faces, _, _ = c_oracle.marching_cubes(vol, 0.0)
        assert len(faces) == g["ntri"][cfg], f"configuration {cfg}: {len(faces)} triangles, classic {g['ntri'][cfg]}"
        want = [int(v) for v in g["patches"][cfg] if v]
        assert _patch_sizes([tuple(f) for f in faces]) == want, f"configuration {cfg}: patches differ"
        assert len(verts) == sum(1 for a in range(3) for p in np.ndindex(2, 2, 2) if p[a] == 0
                                 and (vol[p] > 0) != (vol[tuple(q + (k == a) for k, q in enumerate(p))] > 0))
        checked += 1
    assert checked >= 100  # 256 configurations minus the ones with an ambiguous face


def _edges(faces):
    e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]]).astype(np.int64)
    return e


def _assert_closed_oriented(verts, faces):
    e = _edges(faces)
    n = len(verts)
    fwd, rev = e[:, 0] * n + e[:, 1], e[:, 1] * n + e[:, 0]
    assert len(np.unique(fwd)) == len(fwd), "a directed edge is used twice (non-manifold or inconsistent winding)"
    assert np.array_equal(np.sort(fwd), np.sort(rev)), "an edge has no opposite partner (hole)"


@pytest.mark.parametrize("shape,seed", [((14, 12, 16), 0), ((7, 21, 9), 1), ((20, 20, 20), 2)])
def test_noise_volume_gives_a_closed_oriented_manifold(shape, seed, c_oracle):
    rng = np.random.default_rng(seed)
    vol = np.ones(tuple(s + 2 for s in shape), np.float32)  # a shell of "outside" closes the surface
    vol[1:-1, 1:-1, 1:-1] = rng.standard_normal(shape).astype(np.float32)
    verts, faces, normals, values = c_oracle.marching_cubes(vol, 0.0)
    _assert_closed_oriented(verts, faces)
    assert faces.min() == 0 and faces.max() == len(verts) - 1
    # every vertex sits on a lattice edge whose end points straddle the level
    frac = verts - np.floor(verts)
    on_axis = (frac > 0).sum(axis=1)
    # (tilings 6.1.2, 7.3, 10.2, 12.2, 13.3, 13.4 add a vertex inside the cube -- Lewiner's "c-vertex")
    assert np.isin(on_axis, (0, 1, 3)).all() and (on_axis == 3).mean() < 0.05
    edge = on_axis <= 1
    verts, on_axis, frac = verts[edge], on_axis[edge], frac[edge]
    lo = np.floor(verts).astype(int)
    axis = np.argmax(frac > 0, axis=1)
    hi = lo.copy()
    hi[np.arange(len(hi)), axis] += (on_axis == 1)
    a, b = vol[lo[:, 0], lo[:, 1], lo[:, 2]], vol[hi[:, 0], hi[:, 1], hi[:, 2]]
    assert (((a > 0) != (b > 0)) | (on_axis == 0)).all()
    assert (values == 0).all()


def _euler(verts, faces):
    return len(verts) - len(_edges(faces)) // 2 + len(faces)


def test_euler_characteristic_and_orientation(c_oracle):
    g = np.mgrid[0:40, 0:36, 0:44].astype(np.float32)
    c = np.array([19.3, 17.6, 21.2], np.float32)[:, None, None, None]
    sphere = np.sqrt(((g - c) ** 2).sum(0)) - 11.4
    v, f, nrm, _ = c_oracle.marching_cubes(sphere.astype(np.float32), 0.0)
    _assert_closed_oriented(v, f)
    assert _euler(v, f) == 2
    p = v[f]
    face_n = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
    assert (np.einsum("ij,ij->i", face_n, p.mean(1) - c[:, 0, 0, 0]) > 0).all()  # wound towards increasing values
    # vertex normals follow the gradient direction the wrapper asks for (descent: towards smaller values = inwards)
    radial = v - c[:, 0, 0, 0]
    cosang = np.einsum("ij,ij->i", nrm, radial) / np.linalg.norm(radial, axis=1)
    assert (np.abs(cosang) > 0.9).all() and len(set(np.sign(cosang))) == 1
    q = np.sqrt((g[0] - 19.5) ** 2 + (g[1] - 17.5) ** 2) - 10.0
    torus = np.sqrt(q ** 2 + (g[2] - 21.5) ** 2) - 4.2
    v, f, _, _ = c_oracle.marching_cubes(torus.astype(np.float32), 0.0)
    _assert_closed_oriented(v, f)
    assert _euler(v, f) == 0


def test_two_touching_blobs_ambiguous_faces_stay_watertight(c_oracle):
    g = np.mgrid[0:30, 0:30, 0:30].astype(np.float32)
    d1 = np.sqrt((g[0] - 10) ** 2 + (g[1] - 10) ** 2 + (g[2] - 10) ** 2) - 7.3
    d2 = np.sqrt((g[0] - 19) ** 2 + (g[1] - 19) ** 2 + (g[2] - 19) ** 2) - 7.3
    v, f, _, _ = c_oracle.marching_cubes(np.minimum(d1, d2).astype(np.float32), 0.0)
    _assert_closed_oriented(v, f)
