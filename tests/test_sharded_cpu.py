"""Host-side logic of the multi-GPU paths, on CPU with the gloo backend (world_size 2 and 3): slab partitioning, the seam
halo exchange, the all-gather of the per-column counts, the grouped variable-length gather and the placement of the
per-slab meshes.  The per-slab meshes are cut out of an oracle mesh here and the two device steps (per-column shifts, row
scatter) are numpy stand-ins written in this file (the CUDA kernels cannot run on this box); the same host functions run
over NCCL with the kernels on the GPU box (tests/test_multi_gpu.py, tests/test_marching_cubes_gpu.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hive_b200 import sharded


def test_slab_ranges_cover_and_align():
    for Z in (1, 4, 7, 64, 258, 2048):
        for w in (1, 2, 3, 4, 8):
            r = sharded.slab_ranges(Z, w)
            assert r[0][0] == 0 and r[-1][1] == Z and len(r) == w
            for (a0, a1), (b0, b1) in zip(r[:-1], r[1:]):
                assert a1 == b0 and a0 <= a1
            assert all(a % 4 == 0 or a == Z for a, _ in r)  # empty trailing slabs when Z is tiny


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_slabs(world):
    """The oracle mesh of a noise volume in the product's order, and for every slab: what it owns (vertex / face ids of
    the whole mesh, in slab-local order) and its per-column counts."""
    from oracle import c_oracle
    rng = np.random.default_rng(17)
    vol = rng.standard_normal((11, 9, 24)).astype(np.float32)
    verts, faces, normals, _, keys, cubes = c_oracle.marching_cubes(vol, 0.0, with_cubes=True)
    X, Y, Z = vol.shape
    ranges = sharded.slab_ranges(Z, world)
    colors = (np.arange(len(verts) * 3).reshape(-1, 3) % 251).astype(np.uint8)
    # the whole-volume reference in the product's order: vertices by key, faces by cube (stable)
    order = c_oracle.product_order(keys)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    forder = np.argsort(cubes, kind="stable")
    want = (verts[order], rank[faces[forder]].astype(np.int32), normals[order], colors[order])
    keys_o, cubes_o = keys[order], cubes[forder]
    vz, vcol, centre = (keys_o // 4) % Z, (keys_o // 4) // Z, (keys_o & 3) == 3
    cz, ccol = cubes_o % Z, cubes_o // Z
    slabs = []
    for z0, z1 in ranges:
        vs = np.nonzero((vz >= z0) & (vz < z1))[0]  # ascending in the whole order == the slab's own order
        fs = np.nonzero((cz >= z0) & (cz < z1))[0]
        counts = np.zeros((3, X * Y), np.int32)
        np.add.at(counts[0], vcol[vs][~centre[vs]], 1)
        np.add.at(counts[1], ccol[fs], 1)
        np.add.at(counts[2], vcol[vs][centre[vs]], 1)
        slabs.append({"vs": vs, "fs": fs, "counts": counts, "vcol": vcol[vs], "centre": centre[vs], "ccol": ccol[fs]})
    return vol, ranges, slabs, want


def _column_shifts(all_counts, r):
    """numpy restatement of hv_mc_column_shifts: (vshift, tshift, cshift) per column and the whole-volume totals."""
    tot = all_counts.sum(axis=0, dtype=np.int64)
    low = all_counts[:r].sum(axis=0, dtype=np.int64)
    own = all_counts[r].astype(np.int64)
    excl = lambda a: np.concatenate([[0], np.cumsum(a)[:-1]])  # noqa: E731
    shifts = [excl(tot[k]) + low[k] - excl(own[k]) for k in range(3)]
    return shifts, tot.sum(axis=1)


def _slab_part(slab, want, all_counts, r):
    """What McPlan.emit_placed produces for slab r: its rows in its own order, faces as whole-mesh indices, and the
    destinations derived from the all-gathered counts alone."""
    (vshift, tshift, cshift), (n_edge, n_tri, n_centre) = _column_shifts(all_counts, r)
    vs, fs, centre = slab["vs"], slab["fs"], slab["centre"]
    n_edge_local = int((~centre).sum())
    local = np.arange(len(vs))
    vert_dst = np.where(centre, n_edge + (local - n_edge_local) + cshift[slab["vcol"]], local + vshift[slab["vcol"]])
    tri_dst = np.arange(len(fs)) + tshift[slab["ccol"]]
    assert np.array_equal(vert_dst, vs) and np.array_equal(tri_dst, fs), "the shifts do not reproduce the whole order"
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a))  # noqa: E731
    return {"verts": t(want[0][vs]), "normals": t(want[2][vs]), "colors": t(want[3][vs]), "faces": t(want[1][fs]),
            "vert_dst": t(vert_dst.astype(np.int32)), "tri_dst": t(tri_dst.astype(np.int32))}, int(n_edge + n_centre), int(n_tri)


def _scatter_rows(src, index, dst):
    dst.numpy()[index.numpy()] = src.numpy()


def _worker(rank, world, port, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        vol, ranges, slabs, want = _oracle_slabs(world)
        z0, z1 = ranges[rank]
        # seam exchange: first plane down, last plane up
        t = torch.from_numpy(vol)
        planes = {"tsdf": (t[:, :, z0].contiguous(), t[:, :, z1 - 1].contiguous())}
        halo = sharded.exchange_halo_planes(dist, rank, world, planes)
        if rank > 0:
            assert torch.equal(halo["tsdf_lo"], t[:, :, z0 - 1])
        else:
            assert "tsdf_lo" not in halo
        if rank < world - 1:
            assert torch.equal(halo["tsdf_hi"], t[:, :, z1])
        else:
            assert "tsdf_hi" not in halo
        # per-column counts of every slab, all-gathered; every rank places its own slab from them
        all_counts = sharded.exchange_column_counts(dist, world, torch.from_numpy(slabs[rank]["counts"]))
        assert np.array_equal(all_counts.numpy(), np.stack([s["counts"] for s in slabs]))
        part, n_verts, n_tris = _slab_part(slabs[rank], want, all_counts.numpy(), rank)
        all_sizes = [(len(s["vs"]), len(s["fs"])) for s in slabs]
        gathered = sharded.gather_placed_parts(dist, rank, world, part, torch.device("cpu"), all_sizes)
        if rank == 0:
            verts, faces, normals, colors = sharded.place_mesh_parts(gathered, n_verts, n_tris, _scatter_rows)
            ok = (np.array_equal(verts.numpy(), want[0]) and np.array_equal(faces.numpy(), want[1]) and
                  np.array_equal(normals.numpy(), want[2]) and np.array_equal(colors.numpy(), want[3]))
            with open(result_path, "w") as f:
                f.write("ok" if ok else "mismatch")
        else:
            assert gathered is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_seam_exchange_gather_and_placement_over_gloo(world, c_oracle, tmp_path):
    result = tmp_path / "result.txt"
    mp.spawn(_worker, args=(world, _free_port(), str(result)), nprocs=world, join=True)
    assert result.read_text() == "ok"


def test_placement_detects_missing_rows(c_oracle):
    _, _, slabs, want = _oracle_slabs(2)
    all_counts = np.stack([s["counts"] for s in slabs])
    parts, n_verts, n_tris = [], 0, 0
    for r in range(2):
        part, n_verts, n_tris = _slab_part(slabs[r], want, all_counts, r)
        parts.append(part)
    parts[1] = {k: v[:-1] if k in ("verts", "normals", "colors", "vert_dst") else v for k, v in parts[1].items()}
    with pytest.raises(AssertionError):
        sharded.place_mesh_parts(parts, n_verts, n_tris, _scatter_rows)
