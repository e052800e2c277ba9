"""Multi-GPU paths over NCCL (needs >= 2 GPUs: `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py`): the
z-slab sharded volume integrates and meshes identically to a single-GPU volume."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    from hive_b200.sharded import SlabShardedTSDFVolume
    from hive_b200.synthetic import Scene
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        scene = Scene("small", "room")
        frames = [(f.rgb, f.depth, f.pose) for f in (scene.frame(i) for i in range(9))] if rank == 0 else None
        vol = SlabShardedTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, world_size=world, rank=rank)
        vol.integrate_batch_host(frames, [0, 1, 2], scene.K)          # host frames -> H2D -> broadcast -> integrate
        sf = vol.upload_frames(frames[3:6] if rank == 0 else None, scene.K)
        vol.integrate_batch_device(sf, [0, 1, 2])                     # device-resident frames -> broadcast -> integrate
        tsdf, color = vol.get_volume()
        weight = vol.get_weight_volume()
        verts, faces, norms, colors = vol.get_mesh()   # no collective: seam planes, counts and rows through peer memory + flags
        paths = [vol.last_mesh_path]
        again = [vol.get_mesh(), vol.get_mesh()]        # both parities of the mailboxes, the shared buffers reused
        vol._mc_caps = (64, 64)                         # work records far too small: every slab sees it in the tails and all go round again
        again.append(vol.get_mesh())
        # the volume changes between calls: seam planes, counts and rows of the next epochs must be the NEW ones (nothing a
        # kernel cached from the inboxes' previous contents may survive) -- two more batches, a mesh after each
        later = []
        for batch in ([6, 7], [8]):
            vol.integrate_batch_host(frames, batch, scene.K)
            later.append(vol.get_mesh())
        vol._mb_off = True                              # NCCL for planes / counts / arrival, rows still stored peer to peer
        final = [vol.get_mesh()]
        paths.append(vol.last_mesh_path)
        vol._p2p_off = True                             # and through the grouped send / receive + scatter
        final.append(vol.get_mesh())
        paths.append(vol.last_mesh_path)
        vol.close()
        if rank == 0:
            same = lambda a, b: all(np.array_equal(x, y) for x, y in zip(a, b))  # noqa: E731
            extra = {f"later{k}_{name}": arr for k, m in enumerate(later) for name, arr in zip(("verts", "faces", "norms", "colors"), m)}
            np.savez(out_path, tsdf=tsdf, color=color, weight=weight, verts=verts, faces=faces, norms=norms, colors=colors,
                     paths=np.array(paths), same=np.array([same((verts, faces, norms, colors), m) for m in again]),
                     final_same=np.array([same(later[-1], m) for m in final]), **extra)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sharded_volume_equals_single_gpu(world, tmp_path):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    from hive_b200.fusion import TSDFVolume
    from hive_b200.synthetic import Scene
    out = str(tmp_path / "sharded.npz")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    scene = Scene("small", "room")
    ref = TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size)
    for i in range(6):
        fr = scene.frame(i)
        ref.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
    tsdf, color = ref.get_volume()
    assert np.array_equal(got["tsdf"], tsdf) and np.array_equal(got["color"], color)
    assert np.array_equal(got["weight"], ref.get_weight_volume())
    verts, faces, norms, colors = ref.get_mesh()
    assert np.array_equal(got["verts"], verts) and np.array_equal(got["faces"], faces)
    assert np.array_equal(got["norms"], norms) and np.array_equal(got["colors"], colors)
    assert list(got["paths"]) == ["peer", "p2p", "gather"], "the peer-to-peer mesh assembly did not run"
    assert got["same"].all(), "a repeated mesh differs from the first"
    assert got["final_same"].all(), "the NCCL-coordinated / gathered mesh differs from the collective-free one"
    for k, upto in enumerate((8, 9)):  # the meshes taken after more frames had been integrated
        for i in range(6 if k == 0 else 8, upto):
            fr = scene.frame(i)
            ref.integrate(fr.rgb, fr.depth, scene.K, fr.pose)
        want = ref.get_mesh()
        for name, w in zip(("verts", "faces", "norms", "colors"), want):
            assert np.array_equal(got[f"later{k}_{name}"], w), f"mesh {k} after further frames: {name} differ"
