"""Frame-partitioned foreground meshing (hive/pipeline.py:486-497): the partition and the
variable-length gather of ``hive_b200.pipeline.process_clip`` over gloo on CPU, with a stand-in for the per-frame CUDA step
(tests/test_depth_to_mesh_gpu.py runs the real one on the GPU box)."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from hive_b200 import pipeline


def test_frame_ranges_are_contiguous_balanced_and_complete():
    for n in (0, 1, 5, 16, 1000, 1001):
        for w in (1, 2, 3, 8):
            r = [pipeline.frame_range(n, w, k) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r[:-1], r[1:]))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def _fake_frame(i):
    """A frame whose 'mesh' depends on its index only (ragged: frame 3 is empty, as when no object passes the rules)."""
    n = 0 if i == 3 else 5 + 3 * i
    return (np.full((2, 2, 3), i, np.uint8), np.full((2, 2), float(i), np.float32), np.zeros((2, 2), np.uint8), np.eye(4)), n


def _load(i):
    return _fake_frame(i)[0]


def _stand_in(rgb, depth, mask_encoded, pose, camera_matrix, dilation_options, filtering_options, disable_coverage_constraint):
    i = int(rgb[0, 0, 0])
    n = _fake_frame(i)[1]
    if n == 0:
        return pipeline.FrameMesh(np.zeros((0, 3)), np.zeros((0, 3)))
    v = np.arange(3 * n, dtype=np.float64).reshape(n, 3) + i
    return pipeline.FrameMesh(v, np.zeros((n - 2, 3)) + i, np.full((1, n, 3), i, np.uint8), v[:, :2] / n, (1,))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_frames, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        calls = []

        def load(i):
            calls.append(i)
            return _load(i)

        out = pipeline.process_clip(n_frames, load, np.eye(3), world_size=world, rank=rank, dist=dist, frame_fn=_stand_in)
        start, stop = pipeline.frame_range(n_frames, world, rank)
        assert calls == list(range(start, stop)), "a rank loads exactly its own contiguous frames"
        local_start, local = pipeline.process_clip(n_frames, _load, np.eye(3), world_size=world, rank=rank, dist=dist,
                                                   gather=False, frame_fn=_stand_in)
        assert local_start == start and len(local) == stop - start
        if rank == 0:
            want = pipeline.process_clip(n_frames, _load, np.eye(3), frame_fn=_stand_in)  # the single-process loop
            ok = len(out) == n_frames and all(
                np.array_equal(a.vertices, b.vertices) and np.array_equal(a.faces, b.faces) and a.is_empty == b.is_empty
                and (a.is_empty or (np.array_equal(a.texture, b.texture) and np.array_equal(a.uv, b.uv)))
                for a, b in zip(out, want))
            with open(result_path, "w") as f:
                f.write("ok" if ok else "mismatch")
        else:
            assert out is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames", [(2, 7), (3, 8), (3, 2)])
def test_process_clip_partition_and_gather_over_gloo(world, n_frames, tmp_path):
    result = tmp_path / "result.txt"
    mp.spawn(_worker, args=(world, _free_port(), n_frames, str(result)), nprocs=world, join=True)
    assert result.read_text() == "ok"
