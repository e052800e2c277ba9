#!/usr/bin/env python
"""Where the time of the seam-placed get_mesh of a z-slab sharded volume goes (device sync after every step).

    HV_MESH_TRACE=1 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/trace_sharded_mesh.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    os.environ["HV_MESH_TRACE"] = "1"
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    from hive_b200 import synthetic
    base = synthetic.CONFIGS["C2"]["dim"]
    dim = (base[0], base[1], base[2] * world)
    frames = bench.make_frames("C2", "room", 24, dim) if rank == 0 else None
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hive_b200.sharded import SlabShardedTSDFVolume
    sc = bench._scene("C2", "room", dim)
    vol = SlabShardedTSDFVolume(sc.vol_bnds.copy(), sc.voxel_size, world_size=world, rank=rank)
    dev = vol.upload_frames(frames, sc.K)
    vol.integrate_batch_device(dev, list(range(dev.num_frames)))
    for rep in range(4):
        out = vol.get_mesh()
        tr = vol.last_mesh_trace
        if rank == 0 and rep >= 2:
            tr = tr + [("place+download (wall)", None)]
            steps = [(b[0], 1e3 * (b[1] - a[1])) for a, b in zip(vol.last_mesh_trace[:-1], vol.last_mesh_trace[1:])]
            print(f"rep {rep}: " + ", ".join(f"{n} {ms:.3f}" for n, ms in steps) + f" | vertices {len(out[0])}", flush=True)
    # and untraced (no per-step syncs): wall clock of the whole call on rank 0, both assembly paths
    import time
    os.environ.pop("HV_MESH_TRACE")
    for label in ("peer", "p2p", "gather"):
        if label == "p2p":
            vol._mb_off = True
        if label == "gather":
            vol._p2p_off = True
        ts = []
        for rep in range(7):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            vol.get_mesh()
            ts.append(1e3 * (time.perf_counter() - t0))
        if rank == 0:
            print(f"untraced get_mesh [{vol.last_mesh_path}]: median {sorted(ts[2:])[len(ts[2:]) // 2]:.3f} ms  {['%.3f' % t for t in ts]}", flush=True)
    vol.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
