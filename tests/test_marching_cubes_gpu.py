"""GPU parity of marching cubes / get_mesh against the oracle (oracle/mc_oracle.c, restating fusion.py:338-348).

Gates (BASELINE.json north_star): triangle topology and counts bit-exact, vertex positions within 1e-4 voxel units
(they are in fact bit-equal), colours bit-exact.  The product numbers vertices by (owner voxel, axis) while the serial
oracle numbers them in first-touch order, so meshes are compared after renumbering both through the canonical edge key.
"""
import numpy as np
import pytest

from hive_b200.synthetic import Scene

pytestmark = pytest.mark.gpu


def _canon(verts, faces, normals, keys):
    from oracle import c_oracle
    order = c_oracle.product_order(keys)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    return verts[order], rank[faces].astype(np.int32), normals[order], keys[order]


def _compare(volume, level, c_oracle):
    from hive_b200 import marching_cubes as mc
    ov, of, on, _, ok = c_oracle.marching_cubes(volume, level, with_keys=True)
    ov, of, on, ok = _canon(ov, of, on, ok)
    gv, gf, gn, gvals = mc.marching_cubes(volume, level)
    assert gv.dtype == np.float32 and gf.dtype == np.int32 and gn.dtype == np.float32
    assert gv.shape == ov.shape, "vertex count differs"
    assert gf.shape == of.shape, "triangle count differs"
    assert np.array_equal(gf, of), "triangle topology differs"
    assert np.abs(gv - ov).max() <= 1e-4
    assert np.array_equal(gv, ov)  # same float64 interpolation, same float32 rounding
    assert np.abs(gn - on).max() <= 1e-6
    assert (gvals == np.float32(level)).all()
    return gv, gf


def _sphere(n, centre, radius):
    g = np.mgrid[0:n[0], 0:n[1], 0:n[2]].astype(np.float32)
    c = np.asarray(centre, np.float32)[:, None, None, None]
    return (np.sqrt(((g - c) ** 2).sum(0)) - np.float32(radius)).astype(np.float32)


def test_sphere_matches_oracle_and_is_closed(c_oracle):
    vol = _sphere((48, 40, 70), (23.3, 19.6, 34.2), 15.1)
    v, f = _compare(vol, 0.0, c_oracle)
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]).astype(np.int64)
    fwd, rev = e[:, 0] * len(v) + e[:, 1], e[:, 1] * len(v) + e[:, 0]
    assert len(np.unique(fwd)) == len(fwd) and np.array_equal(np.sort(fwd), np.sort(rev))  # closed 2-manifold
    assert len(v) - len(e) // 2 + len(f) == 2  # Euler characteristic of a sphere
    p = v[f]
    outward = np.einsum("ij,ij->i", np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]), p.mean(1) - [23.3, 19.6, 34.2])
    assert (outward > 0).all()  # right-hand normals towards increasing values


@pytest.mark.parametrize("shape", [(24, 20, 28), (9, 33, 65), (5, 4, 2), (2, 2, 2), (3, 3, 97)])
def test_noise_volume_exercises_every_ambiguous_case(shape, c_oracle):
    rng = np.random.default_rng(sum(shape))
    _compare(rng.standard_normal(shape).astype(np.float32), 0.0, c_oracle)


def test_every_reachable_lewiner_subcase_gives_the_app_c_triangle_count(c_oracle):
    """Crafted corner values for every (case, sub-case) of Lewiner's decision tree that values can reach, laid out as
    2x2x2 blocks along z: per-cube triangle counts equal SURVEY.md App. C's inventory (the only reference-derived
    numbers on this box), and the GPU mesh equals the oracle's, triangle for triangle."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import gen_mc_tables as gen
    from test_mc_tables import APP_C_TRIANGLES, _random_cubes
    per_label = {}
    for v in _random_cubes(60000, 11):
        label = gen.decide(v)[3]
        if label != "0" and len(per_label.setdefault(label, [])) < 6:
            per_label[label].append(v.astype(np.float32))
    assert set(APP_C_TRIANGLES) - set(per_label) <= {"12.1.2", "13.5.2", "7.4.2"}  # see tests/test_mc_tables.py
    blocks = [(label, v) for label in sorted(per_label) for v in per_label[label]]
    vol = np.zeros((2, 2, 2 * len(blocks)), np.float32)
    for b, (label, v) in enumerate(blocks):
        for i in range(8):
            o = gen.CORNERS[i]
            vol[o[0], o[1], 2 * b + o[2]] = v[i]
    _compare(vol, 0.0, c_oracle)
    _, faces, _, _, _, cubes = c_oracle.marching_cubes(vol, 0.0, with_cubes=True)
    per_cube = np.bincount(cubes, minlength=2 * len(blocks))
    for b, (label, v) in enumerate(blocks):
        # float32 storage can move a value across a decision boundary only if the float64 decision was a near-tie
        if gen.decide(v.astype(np.float64))[3] == label:
            assert per_cube[2 * b] == APP_C_TRIANGLES[label], (label, v)


def test_nonzero_level_and_exact_zeros(c_oracle):
    rng = np.random.default_rng(3)
    vol = rng.integers(-3, 4, size=(17, 19, 40)).astype(np.float32)  # many voxels exactly on the level
    _compare(vol, 0.0, c_oracle)
    _compare(vol, 0.5, c_oracle)
    _compare(rng.random((20, 20, 20)).astype(np.float32), 0.37, c_oracle)


def test_errors_match_skimage_contract(c_oracle):
    from hive_b200 import marching_cubes as mc
    ones = np.ones((8, 8, 8), np.float32)
    with pytest.raises(ValueError):  # level below the data range: what an untouched TSDF volume gives (experiments.py:167)
        mc.marching_cubes(ones, 0.0)
    with pytest.raises(ValueError):
        c_oracle.marching_cubes(ones, 0.0)
    with pytest.raises(RuntimeError):  # level inside the range but no voxel strictly above it
        mc.marching_cubes(np.zeros((8, 8, 8), np.float32), 0.0)
    with pytest.raises(RuntimeError):
        c_oracle.marching_cubes(np.zeros((8, 8, 8), np.float32), 0.0)
    with pytest.raises(ValueError):
        mc.marching_cubes(np.zeros((8, 8), np.float32), 0.0)


@pytest.mark.parametrize("config,variant,frames", [("tiny", "room", 6), ("small", "room", 8), ("small", "dense", 3)])
def test_get_mesh_matches_oracle(config, variant, frames, c_oracle):
    from hive_b200.fusion import TSDFVolume
    scene = Scene(config, variant)
    gpu = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    orc = c_oracle.COracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for i in range(frames):
        fr = scene.frame(i)
        gpu.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        orc.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
    tsdf, color = gpu.get_volume()
    assert np.array_equal(color, orc._color_vol_cpu)
    # oracle mesh of the GPU's own planes (isolates marching cubes from the 1e-5 TSDF tolerance)
    ov, of, on, _, ok = c_oracle.marching_cubes(tsdf, 0.0, with_keys=True)
    ov, of, on, ok = _canon(ov, of, on, ok)
    from oracle import fusion_oracle as fo
    ow, oc = fo.mesh_postprocess(ov, color, gpu._voxel_size, gpu._vol_origin)
    verts, faces, norms, colors = gpu.get_mesh()
    assert verts.dtype == np.float32 and faces.dtype == np.int32 and norms.dtype == np.float32 and colors.dtype == np.uint8
    assert verts.shape == ow.shape and len(verts) > 100
    assert np.array_equal(faces, of)
    assert np.array_equal(verts, ow)
    assert np.array_equal(colors, oc)
    assert np.abs(norms - on).max() <= 1e-6
    # the object stays usable after get_mesh (the reference pops its CUDA context, fusion.py:349)
    pc = gpu.get_point_cloud()
    assert pc.shape == (len(verts), 6) and np.array_equal(pc[:, :3], verts)


def test_get_mesh_on_untouched_volume_raises_value_error():
    from hive_b200.fusion import TSDFVolume
    scene = Scene("tiny", "room")
    vol = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    with pytest.raises(ValueError):
        vol.get_mesh()


@pytest.mark.parametrize("n_slabs", [2, 3])
def test_slab_halo_stitch_equals_whole_volume(n_slabs):
    """z-slab shards emulated on one GPU: per-slab extraction with halo planes, placed by the per-column counts of all
    slabs == whole-volume mesh, bit for bit."""
    from hive_b200 import marching_cubes as mc
    from hive_b200 import sharded
    from hive_b200.fusion import TSDFVolume
    scene = Scene("small", "room")
    whole = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    Zg = int(whole._vol_dim[2])
    ranges = sharded.slab_ranges(Zg, n_slabs)
    slabs = [TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, z_range=r) for r in ranges]
    for i in range(5):
        fr = scene.frame(i)
        for v in [whole] + slabs:
            v.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
    # slab planes are the corresponding slices of the whole volume, bit for bit
    tsdf, color = whole.get_volume()
    for s, (z0, z1) in zip(slabs, ranges):
        st, sc = s.get_volume()
        assert np.array_equal(st, tsdf[:, :, z0:z1]) and np.array_equal(sc, color[:, :, z0:z1])
    import torch
    plans, sizes, counts = [], [], []
    for k, (s, (z0, z1)) in enumerate(zip(slabs, ranges)):
        Z = z1 - z0
        halo = {}
        if k > 0:
            lo = slabs[k - 1]
            zl = int(lo._local_dim[2]) - 1
            halo["tsdf_lo"] = lo._tsdf_vol_gpu[:, :, zl].contiguous()
            halo["color_lo"] = lo._color_vol_gpu[:, :, zl].contiguous()
        if k < n_slabs - 1:
            hi = slabs[k + 1]
            halo["tsdf_hi"] = hi._tsdf_vol_gpu[:, :, 0].contiguous()
            halo["color_hi"] = hi._color_vol_gpu[:, :, 0].contiguous()
        X, Y = int(s._local_dim[0]), int(s._local_dim[1])
        plan = mc.McPlan(s._tsdf_vol_gpu, s._color_vol_gpu, (X, Y, Z), s._pitch, halo=halo, z_global_offset=z0,
                         z_global_dim=Zg)
        plan.count(64, 64)  # far too small on purpose: the records do not fit and the count is repeated, exactly sized
        nv, nt, _, _ = plan.read_totals()
        plan.ensure_records(nv, nt)
        plans.append(plan)
        sizes.append((nv, nt))
        counts.append(plan.column_counts())
    all_counts = torch.stack(counts)  # what the all-gather of the NCCL path produces
    parts = []
    for k, (plan, (nv, nt), s) in enumerate(zip(plans, sizes, slabs)):
        shifts, g_tot = plan.column_shifts(all_counts, k)
        n_edge, n_tri, n_centre = (int(v) for v in g_tot[:3].tolist())
        tails = all_counts[:, -plan.SLAB_TAIL:].cpu().numpy()  # every slab's (min key, max key, vertices, triangles, edge vertices)
        assert [tuple(int(v) for v in t[2:4]) for t in tails] == sizes and int(tails[:, 4].sum()) == n_edge
        parts.append(plan.emit_placed(nv, nt, s._voxel_size, s._vol_origin, shifts, n_edge))
    assert sum(nv for nv, _ in sizes) == n_edge + n_centre and sum(nt for _, nt in sizes) == n_tri
    verts, faces, normals, colors = (t.cpu().numpy() for t in
                                     sharded.place_mesh_parts(parts, n_edge + n_centre, n_tri, mc.scatter_rows))
    wv, wf, wn, wc = whole.get_mesh()
    assert len(wv) > 1000
    assert np.array_equal(verts, wv) and np.array_equal(faces, wf)
    assert np.array_equal(normals, wn) and np.array_equal(colors, wc)
    # direct placement (hv_mc_emit_direct, the peer-to-peer path of sharded.get_mesh): every slab stores its rows at their
    # final positions in ONE buffer laid out like the one rank 0 shares with the other ranks
    from hive_b200 import peer
    n_v, n_t = n_edge + n_centre, n_tri
    offs, total = peer.mesh_layout(n_v, n_t)
    buf = torch.full((total,), 0xCD, dtype=torch.uint8, device=whole.device)
    assert buf.data_ptr() % 256 == 0
    for k, (plan, (nv, nt), s) in enumerate(zip(plans, sizes, slabs)):
        shifts, _ = plan.column_shifts(all_counts, k)
        plan.emit_direct(nv, nt, s._voxel_size, s._vol_origin, shifts, n_edge, n_v, n_t,
                         {name: buf.data_ptr() + o for name, o in offs.items()})
    torch.cuda.synchronize()
    host = buf.cpu().numpy()
    assert np.array_equal(host[offs["verts"]:offs["verts"] + 12 * n_v].view(np.float32).reshape(-1, 3), wv)
    assert np.array_equal(host[offs["normals"]:offs["normals"] + 12 * n_v].view(np.float32).reshape(-1, 3), wn)
    assert np.array_equal(host[offs["colors"]:offs["colors"] + 3 * n_v].reshape(-1, 3), wc)
    assert np.array_equal(host[offs["faces"]:offs["faces"] + 12 * n_t].view(np.int32).reshape(-1, 3), wf)
    # nothing outside the arrays was touched
    used = np.zeros(total, dtype=bool)
    for name, nbytes in (("verts", 12 * n_v), ("normals", 12 * n_v), ("colors", 3 * n_v), ("faces", 12 * n_t)):
        used[offs[name]:offs[name] + nbytes] = True
    assert np.all(host[~used] == 0xCD)


def test_one_call_extract_and_capacity_fallback():
    """hv_mc_extract (count + emit in one library call): with room to spare it equals get_mesh; with a capacity far too
    small it reports the sizes and the plan falls back to an exactly sized count + emit -- same mesh."""
    from hive_b200 import marching_cubes as mc
    from hive_b200.fusion import TSDFVolume
    scene = Scene("small", "room")
    vol = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for i in range(4):
        fr = scene.frame(i)
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
    want = vol.get_mesh()
    X, Y, Z = (int(d) for d in vol._local_dim)
    for cap_v, cap_t in ((4 * len(want[0]), 4 * len(want[1])), (100, 100)):
        plan = mc.McPlan(vol._tsdf_vol_gpu, vol._color_vol_gpu, (X, Y, Z), vol._pitch)
        mesh = plan.extract(vol._voxel_size, vol._vol_origin, cap_v, cap_t)
        got = mesh.to_host()
        for a, b in zip(got, want):
            assert a.shape == b.shape and np.array_equal(a, b)
        assert mesh.vmin <= 0.0 <= mesh.vmax
