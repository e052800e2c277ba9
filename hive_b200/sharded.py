"""z-slab sharding of one TSDF volume over the GPUs of a B200 box (BASELINE.json north_star (a)).

The reference is single-GPU (device 0 hard-coded, fusion.py:65); this layer is new design, not a port.  Voxels are
independent given the frame, so integration needs no reduction: the global ``[X][Y][Z]`` volume is cut along z into
``world_size`` contiguous slabs, one dense C-order array per GPU.  Each rank integrates every frame into its slab; the
slab's slice of the global z world-coordinate table makes the result identical to the unsharded volume.  The only
data-path collective is the NCCL broadcast of each frame batch (packed texels) from rank 0, double-buffered so the
broadcast of batch s+1 overlaps the integration of batch s; camera poses are tiny and broadcast once, up front.

One process per GPU (``torch.distributed``); ``world_size == 1`` degenerates to a plain ``TSDFVolume``.
"""
from __future__ import annotations

import numpy as np

from .fusion import DeviceFrames, TSDFVolume


def slab_ranges(Z: int, world_size: int):
    """Contiguous z ranges, as even as possible, boundaries on multiples of 4 voxels (float4 alignment of the seam)."""
    bounds = [min(Z, (Z * r // world_size + 3) // 4 * 4) for r in range(world_size)] + [Z]
    bounds[0] = 0
    return [(bounds[r], bounds[r + 1]) for r in range(world_size)]


class ShardedFrames:
    """Frames owned by rank 0 (device texels) plus camera data replicated on every rank."""

    def __init__(self, dev: DeviceFrames | None, world2cam, intr, H, W, num_frames):
        self.dev, self.world2cam, self.intr, self.H, self.W, self.num_frames = dev, world2cam, intr, H, W, num_frames


class SlabShardedTSDFVolume:
    def __init__(self, vol_bnds, voxel_size, world_size=1, rank=0, device=None):
        import torch
        self._torch = torch
        self.world_size, self.rank = int(world_size), int(rank)
        vol_bnds = np.array(vol_bnds, dtype=np.float64)
        gdim = np.ceil((vol_bnds[:, 1] - vol_bnds[:, 0]) / float(voxel_size)).astype(int)
        self.global_dim = tuple(int(d) for d in gdim)
        self.ranges = slab_ranges(self.global_dim[2], self.world_size)
        z0, z1 = self.ranges[self.rank]
        assert z1 > z0, "more ranks than z-slabs"
        self.local = TSDFVolume(vol_bnds, voxel_size, device=device, z_range=(z0, z1) if self.world_size > 1 else None)
        self._bcast = [None, None]
        self._step = 0
        if self.world_size > 1:
            import torch.distributed as dist
            self._dist = dist

    # ------------------------------------------------------------------ update counter
    def count_updates(self, flag: bool):
        self.local.count_updates = bool(flag)

    def reset_update_counter(self):
        self.local.reset_update_counter()

    def num_updated_local(self) -> int:
        return self.local.num_updated()

    def read_update_counter(self) -> int:
        return self.local.num_updated()

    # ------------------------------------------------------------------ frames
    def upload_frames(self, frames, cam_intr) -> ShardedFrames:
        """``frames``: list of (color_im, depth_im, cam_pose) on rank 0 (None elsewhere)."""
        if self.world_size == 1:
            dev = self.local.upload_frames(frames, cam_intr)
            return ShardedFrames(dev, dev.world2cam, dev.intr, dev.H, dev.W, len(dev))
        meta = [None]
        dev = None
        if self.rank == 0:
            dev = self.local.upload_frames(frames, cam_intr)
            meta = [(dev.world2cam, dev.intr, dev.H, dev.W, len(dev))]
        self._dist.broadcast_object_list(meta, src=0)
        w2c, intr, H, W, n = meta[0]
        return ShardedFrames(dev, w2c, intr, H, W, n)

    def _batch_buffer(self, slot, B, H, W):
        torch = self._torch
        buf = self._bcast[slot]
        if buf is None or buf.shape != (B, H * W):
            buf = torch.empty((B, H * W), dtype=torch.int64, device=self.local.device)
            self._bcast[slot] = buf
        return buf

    def integrate_batch_device(self, sf: ShardedFrames, idx, obs_weight=1.0):
        """Integrate frames ``idx`` (already resident in rank 0's HBM) into every slab."""
        if self.world_size == 1:
            for i in idx:
                self.local.integrate_device(sf.dev, i, obs_weight)
            return
        torch = self._torch
        slot = self._step & 1
        self._step += 1
        buf = self._batch_buffer(slot, len(idx), sf.H, sf.W)
        if self.rank == 0:
            torch.index_select(sf.dev.texels, 0, torch.as_tensor(idx, device=buf.device), out=buf)
        self._dist.broadcast(buf, src=0)  # NCCL over NVLink; stream-ordered after the producers/consumers of `buf`
        frames = DeviceFrames(buf, [sf.world2cam[i] for i in idx], sf.intr, sf.H, sf.W)
        for j in range(len(idx)):
            self.local.integrate_device(frames, j, obs_weight)

    def integrate_batch_host(self, frames, idx, cam_intr, obs_weight=1.0):
        """End-to-end path: host numpy frames (rank 0) -> H2D -> (broadcast) -> integrate."""
        if self.world_size == 1:
            for i in idx:
                color_im, depth_im, pose = frames[i]
                self.local.integrate(color_im, depth_im, cam_intr, pose, obs_weight)
            return
        torch = self._torch
        H, W = (int(s) for s in (frames[idx[0]][1].shape if self.rank == 0 else (0, 0)))
        shape = torch.tensor([H, W], dtype=torch.int64, device=self.local.device)
        self._dist.broadcast(shape, src=0)
        H, W = (int(v) for v in shape.tolist())
        slot = self._step & 1
        self._step += 1
        buf = self._batch_buffer(slot, len(idx), H, W)
        cams = torch.empty((len(idx), 16), dtype=torch.float64, device=self.local.device)
        if self.rank == 0:
            host_cams = np.zeros((len(idx), 16))
            for j, i in enumerate(idx):
                color_im, depth_im, pose = frames[i]
                intr, w2c = self.local._camera(cam_intr, pose)
                host_cams[j, :12] = w2c.reshape(-1)
                host_cams[j, 12:] = intr
                self.local.upload_into(color_im, depth_im, buf[j])
            cams.copy_(torch.from_numpy(host_cams))
        self._dist.broadcast(buf, src=0)
        self._dist.broadcast(cams, src=0)
        hc = cams.cpu().numpy()
        frames_dev = DeviceFrames(buf, [np.ascontiguousarray(hc[j, :12]) for j in range(len(idx))],
                                  hc[0, 12:].astype(np.float32), H, W)
        for j in range(len(idx)):
            self.local.integrate_device(frames_dev, j, obs_weight)

    # ------------------------------------------------------------------ read-back
    def get_volume(self):
        """(tsdf, color) of the *global* volume on rank 0 (None elsewhere): slabs concatenated along z."""
        tsdf, color = self.local.get_volume()
        if self.world_size == 1:
            return tsdf, color
        parts = [None] * self.world_size if self.rank == 0 else None
        self._dist.gather_object((tsdf, color), parts, dst=0)
        if self.rank != 0:
            return None, None
        return (np.concatenate([p[0] for p in parts], axis=2), np.concatenate([p[1] for p in parts], axis=2))

    def get_weight_volume(self):
        w = self.local.get_weight_volume()
        if self.world_size == 1:
            return w
        parts = [None] * self.world_size if self.rank == 0 else None
        self._dist.gather_object(w, parts, dst=0)
        return np.concatenate(parts, axis=2) if self.rank == 0 else None
