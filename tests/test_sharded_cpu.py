"""Host-side logic of the multi-GPU paths, on CPU with the gloo backend (world_size 2): slab partitioning, the seam
halo exchange, the variable-length gather and the key-based stitch of per-slab meshes.  The per-slab meshes are cut out
of an oracle mesh here (the CUDA kernels cannot run on this box); the same functions run over NCCL on the GPU box
(tests/test_multi_gpu.py)."""
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


def _oracle_parts(world):
    from oracle import c_oracle
    rng = np.random.default_rng(17)
    vol = rng.standard_normal((11, 9, 24)).astype(np.float32)
    verts, faces, normals, _, keys, cubes = c_oracle.marching_cubes(vol, 0.0, with_cubes=True)
    Z = vol.shape[2]
    ranges = sharded.slab_ranges(Z, world)
    colors = (np.arange(len(verts) * 3).reshape(-1, 3) % 251).astype(np.uint8)
    vz = (keys // 4) % Z
    cz = cubes % Z
    parts = []
    for z0, z1 in ranges:
        vs = (vz >= z0) & (vz < z1)
        fs = (cz >= z0) & (cz < z1)
        parts.append({"verts": verts[vs], "normals": normals[vs], "colors": colors[vs], "vert_keys": keys[vs],
                      "face_keys": keys[faces[fs]], "face_cubes": cubes[fs]})
    # the whole-volume reference in the product's order: vertices by key, faces by cube (stable)
    order = c_oracle.product_order(keys)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    forder = np.argsort(cubes, kind="stable")
    want = (verts[order], rank[faces[forder]].astype(np.int32), normals[order], colors[order])
    return vol, ranges, parts, want


def _worker(rank, world, port, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        vol, ranges, parts, want = _oracle_parts(world)
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
        part = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in parts[rank].items()}
        gathered = sharded.gather_mesh_parts(dist, rank, world, part, torch.device("cpu"))
        if rank == 0:
            verts, faces, normals, colors = sharded.stitch_mesh_parts(gathered)
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
def test_seam_exchange_gather_and_stitch_over_gloo(world, c_oracle, tmp_path):
    result = tmp_path / "result.txt"
    mp.spawn(_worker, args=(world, _free_port(), str(result)), nprocs=world, join=True)
    assert result.read_text() == "ok"


def test_stitch_detects_duplicate_seam_vertices(c_oracle):
    _, _, parts, _ = _oracle_parts(2)
    tp = [{k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in p.items()} for p in parts]
    tp[1]["vert_keys"][0] = tp[0]["vert_keys"][0]
    with pytest.raises(AssertionError):
        sharded.stitch_mesh_parts(tp)
