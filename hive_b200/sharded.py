"""z-slab sharding of one TSDF volume over the GPUs of a B200 box (BASELINE.json north_star (a)).

The reference is single-GPU (device 0 hard-coded, fusion.py:65); this layer is new design, not a port.  Voxels are
independent given the frame, so integration needs no reduction: the global ``[X][Y][Z]`` volume is cut along z into
``world_size`` contiguous slabs, one dense C-order array per GPU.  Each rank integrates every frame into its slab; the
slab's slice of the global z world-coordinate table makes the result identical to the unsharded volume.  The only
data-path collective of integration is the NCCL broadcast of each frame batch (packed texels) from rank 0,
double-buffered so the broadcast of batch s+1 overlaps the integration of batch s; camera poses are tiny and broadcast
once, up front.  Mesh extraction exchanges one boundary plane per seam and one table of per-column counts; every slab then
stores its rows at their final positions in the whole mesh, which lives in rank 0's HBM and is mapped into every rank
(CUDA IPC over NVLink, ``hive_b200/peer.py``) -- or, where the ranks cannot share memory, sends them to rank 0 (``get_mesh``).

One process per GPU (``torch.distributed``); ``world_size == 1`` degenerates to a plain ``TSDFVolume``.
"""
from __future__ import annotations

import os
import time

import numpy as np

from . import _xfer, peer
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
        self._consumed = [None, None]  # event: the integrate kernels that read slot i's texels have finished
        self._step = 0
        self._counter_ring = None
        self._frame_hw = None
        if self.world_size > 1:
            import torch.distributed as dist
            self._dist = dist
            with torch.cuda.device(self.local.device):
                self._copy_stream = torch.cuda.Stream(device=self.local.device)

    # ------------------------------------------------------------------ update counter
    def count_updates(self, flag: bool):
        self.local.count_updates = bool(flag)

    def reset_update_counter(self):
        self.local.reset_update_counter()

    def num_updated_local(self) -> int:
        return self.local.num_updated()

    def read_update_counter(self) -> int:
        return self.local.num_updated()

    def read_update_counter_pipelined(self) -> int:
        """Device->host read of the update counter without draining the stream: the counter is snapshotted into a
        pinned ring by an async copy ordered after the kernels enqueued so far, and the value returned is the snapshot
        taken by the *previous* call (0 on the first), whose copy has long finished.  Every call still moves its 8 bytes
        to the host; ``read_update_counter`` gives the synchronous, current value."""
        torch = self._torch
        if self._counter_ring is None:
            self._counter_ring = {"host": [torch.zeros(1, dtype=torch.int64).pin_memory() for _ in range(2)],
                                  "event": [None, None], "i": 0}
        ring = self._counter_ring
        i = ring["i"]
        ring["i"] = i ^ 1
        with torch.cuda.device(self.local.device):
            ring["host"][i].copy_(self.local._updated, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            ring["event"][i] = ev
        prev = ring["event"][i ^ 1]
        if prev is None:
            return 0
        prev.synchronize()
        return int(ring["host"][i ^ 1].item())

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
        dev = self.local.device
        slot = self._step & 1
        self._step += 1
        idx = [int(i) for i in idx]
        contiguous = all(b == a + 1 for a, b in zip(idx[:-1], idx[1:]))
        with torch.cuda.device(dev):
            compute = torch.cuda.current_stream()
            copy = self._copy_stream
            # The broadcast of this batch runs on the side stream, behind the integration that last used the slot (two
            # batches ago) and therefore next to the integration of the previous batch, which is still on the compute stream.
            if self._consumed[slot] is not None:
                copy.wait_event(self._consumed[slot])
            else:
                copy.wait_stream(compute)
            with torch.cuda.stream(copy):
                if self.rank == 0 and contiguous:
                    # rank 0 sends straight out of its resident frames and integrates from them: no staging copy
                    buf = sf.dev.texels[idx[0]:idx[0] + len(idx)]
                else:
                    buf = self._batch_buffer(slot, len(idx), sf.H, sf.W)
                    if self.rank == 0:
                        torch.index_select(sf.dev.texels, 0, torch.as_tensor(idx, device=buf.device), out=buf)
                self._dist.broadcast(buf, src=0)  # NCCL over NVLink
                ready = torch.cuda.Event()
                ready.record(copy)
            compute.wait_event(ready)
            frames = DeviceFrames(buf, [sf.world2cam[i] for i in idx], sf.intr, sf.H, sf.W)
            for j in range(len(idx)):
                self.local.integrate_device(frames, j, obs_weight)
            done = torch.cuda.Event()
            done.record(compute)
            self._consumed[slot] = done  # the side stream must not refill the slot before this

    def integrate_batch_host(self, frames, idx, cam_intr, obs_weight=1.0):
        """End-to-end path: host numpy frames (rank 0) -> H2D -> (broadcast) -> integrate."""
        if self.world_size == 1:
            for i in idx:
                color_im, depth_im, pose = frames[i]
                self.local.integrate(color_im, depth_im, cam_intr, pose, obs_weight)
            return
        torch = self._torch
        dev = self.local.device
        slot = self._step & 1
        self._step += 1
        with torch.cuda.device(dev):
            compute = torch.cuda.current_stream()
            copy = self._copy_stream
            # Upload + broadcast run on a side stream so that they overlap the integration of the previous batch (the
            # caller is not blocked on the compute stream); the slot is reused every other batch.
            if self._consumed[slot] is not None:
                copy.wait_event(self._consumed[slot])
            else:
                copy.wait_stream(compute)
            with torch.cuda.stream(copy):
                H, W = (int(v) for v in (frames[idx[0]][1].shape if self.rank == 0 else (0, 0)))
                meta = torch.tensor([H, W], dtype=torch.int64, device=dev)
                if self._frame_hw is None:  # frame size is fixed for the life of the volume: broadcast it once
                    self._dist.broadcast(meta, src=0)
                    self._frame_hw = tuple(int(v) for v in meta.tolist())
                H, W = self._frame_hw
                buf = self._batch_buffer(slot, len(idx), H, W)
                cams = torch.empty((len(idx), 16), dtype=torch.float64, device=dev)
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
                hc = cams.cpu().numpy()  # waits for the side stream only
                ready = torch.cuda.Event()
                ready.record(copy)
            compute.wait_event(ready)
            frames_dev = DeviceFrames(buf, [np.ascontiguousarray(hc[j, :12]) for j in range(len(idx))],
                                      hc[0, 12:].astype(np.float32), H, W)
            for j in range(len(idx)):
                self.local.integrate_device(frames_dev, j, obs_weight)
            done = torch.cuda.Event()
            done.record(compute)
            self._consumed[slot] = done

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

    # ------------------------------------------------------------------ mesh extraction (seam-stitched marching cubes)
    def _boundary_planes(self):
        """(first, last) z-planes of this slab's tsdf and colour volumes as (X, Y) views."""
        loc = self.local
        Z = int(loc._local_dim[2])
        t, c = loc._tsdf_vol_gpu, loc._color_vol_gpu
        return t[:, :, 0], t[:, :, Z - 1], c[:, :, 0], c[:, :, Z - 1]  # strided views: torch.stack packs them in one pass each

    def get_mesh(self):
        """``TSDFVolume.get_mesh`` of the global volume on rank 0 (``None`` x4 elsewhere).

        Each slab owns the cubes whose base voxel it holds; the planes either side of a seam are exchanged once
        (one tsdf + one colour plane per direction).  The whole-volume mesh is ordered column by column, so the slabs
        interleave: every slab counts what it owns per column, the counts are all-gathered, and every slab works out by
        itself where its vertices and triangles go in the whole mesh and writes its faces with whole-mesh indices (seam
        vertices, owned by the slab above, included).  The emit kernels then store every row at its final position in
        rank 0's HBM over NVLink (``hv_mc_emit_direct``; one small all-reduce tells rank 0 that all rows have arrived);
        where the ranks cannot share memory (gloo, HV_MESH_P2P=0), rank 0 receives the slab meshes in one grouped exchange
        and drops every row into place.  The result is identical -- vertex and face order included -- to the mesh of the
        unsharded volume; nothing is sorted or searched.
        """
        from . import marching_cubes as mc
        if self.world_size == 1:
            return self.local.get_mesh()
        dist, loc, torch = self._dist, self.local, self._torch
        z0, z1 = self.ranges[self.rank]
        Zg = self.global_dim[2]
        trace = [] if os.environ.get("HV_MESH_TRACE") else None  # (step, ms) with a device sync per step: tools/trace_sharded_mesh.py

        def mark(step):
            if trace is not None:
                torch.cuda.synchronize()
                trace.append((step, time.perf_counter()))

        mark("start")
        X, Y, Z = (int(d) for d in loc._local_dim)
        if self._peer_mailboxes(X * Y) is not None:
            return self._get_mesh_peer(mark, trace)
        first_t, last_t, first_c, last_c = self._boundary_planes()
        mark("boundary planes")
        # the halo planes and the extraction plan (scratch, work records) live as long as the volume: a repeated get_mesh
        # allocates nothing.  tsdf and colour plane travel as one message per neighbour.
        packed = exchange_halo_planes(dist, self.rank, self.world_size,
                                      {"tc": (torch.stack((first_t, first_c)), torch.stack((last_t, last_c)))},
                                      out=getattr(self, "_halo_packed", None))
        if getattr(self, "_halo_packed", None) is None:
            self._halo_packed = packed
            self._halo = {}
            for side in ("lo", "hi"):
                if f"tc_{side}" in packed:
                    self._halo[f"tsdf_{side}"], self._halo[f"color_{side}"] = packed[f"tc_{side}"][0], packed[f"tc_{side}"][1]
        mark("halo exchange")
        X, Y, Z = (int(d) for d in loc._local_dim)
        plan = getattr(self, "_mc_plan", None)
        if plan is None or plan.tsdf is not loc._tsdf_vol_gpu:
            plan = self._mc_plan = mc.McPlan(loc._tsdf_vol_gpu, loc._color_vol_gpu, (X, Y, Z), loc._pitch, halo=self._halo,
                                             z_global_offset=z0, z_global_dim=Zg)
        plan.count(*getattr(self, "_mc_caps", (None, None)))
        n_verts, n_tris, _, _ = plan.read_totals()
        plan.ensure_records(n_verts, n_tris)
        mark("count + totals")
        self._mc_caps = (n_verts + n_verts // 2 + 1024, n_tris + n_tris // 2 + 1024)  # next time: records sized by this surface
        # ONE collective tells every slab everything about every other: per-column counts + (min, max, sizes)
        counts = plan.column_counts()
        mark("column counts")
        all_counts = exchange_column_counts(dist, self.world_size, counts)
        mark("all-gather")
        shifts, _ = plan.column_shifts(all_counts, self.rank)
        mark("shifts")
        # (world, 5): min key, max key, vertices, triangles, edge vertices of every slab -- the one host read of the exchange
        tails = self._read_tails(all_counts[:, -plan.SLAB_TAIL:])
        vmin = min(mc._key_to_float(int(k)) for k in tails[:, 0])
        vmax = max(mc._key_to_float(int(k)) for k in tails[:, 1])
        n_verts_g, n_tris_g, n_edge_g = (int(v) for v in tails[:, 2:5].sum(axis=0))
        mc.check_surface(0.0, vmin, vmax, n_verts_g)
        mark("tails read")
        buf = self._shared_mesh_buffer(n_verts_g, n_tris_g)
        if buf is not None:
            # peer-to-peer assembly: every slab's emit kernel stores its rows at their final positions in rank 0's HBM
            # (over NVLink from the other GPUs).  The all-reduce is the arrival signal: it is stream-ordered behind the emit
            # kernel on every rank, so when rank 0's completes every slab's rows have landed.
            offs, _ = peer.mesh_layout(n_verts_g, n_tris_g)
            plan.emit_direct(n_verts, n_tris, loc._voxel_size, loc._vol_origin, shifts, n_edge_g, n_verts_g, n_tris_g,
                             {name: buf.base + o for name, o in offs.items()})
            mark("emit into rank 0")
            dist.all_reduce(self._mesh_done)
            mark("arrival")
            self.last_mesh_trace, self.last_mesh_path = trace, "p2p"
            if self.rank != 0:
                return None, None, None, None
            out = self._download_shared_mesh(buf, offs, n_verts_g, n_tris_g)
            mark("download")
            return out
        self.last_mesh_path = "gather"
        part = plan.emit_placed(n_verts, n_tris, loc._voxel_size, loc._vol_origin, shifts, n_edge_g)
        mark("emit")
        parts = gather_placed_parts(dist, self.rank, self.world_size, part, loc.device, tails[:, 2:4])
        mark("gather")
        self.last_mesh_trace = trace
        if self.rank != 0:
            return None, None, None, None
        verts, faces, normals, colors = place_mesh_parts(parts, n_verts_g, n_tris_g, mc.scatter_rows)
        mark("place")
        out = _xfer.to_host(verts, faces, normals, colors)
        mark("download")
        return out


    def _read_tails(self, tails_dev):
        torch = self._torch
        if not tails_dev.is_cuda:
            return tails_dev.numpy()
        host = getattr(self, "_tails_host", None)
        if host is None or host.shape != tails_dev.shape:
            host = self._tails_host = torch.empty(tails_dev.shape, dtype=tails_dev.dtype, pin_memory=True)
        host.copy_(tails_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return host.numpy().copy()

    # -- the slabs talk through each other's memory: seam planes, per-column counts, flags (hive_b200/peer.py) ------------------
    _FLAG_WORDS = 64  # per mailbox: [0] lo halo arrived, [1] hi halo arrived, [8 + w] counts of slab w ready, [24 + w] rows of slab w stored

    def _peer_mailboxes(self, n_cols):
        """Every rank's mailbox mapped into every rank, or None when that is not possible (decided collectively, once)."""
        dist, torch = self._dist, self._torch
        if getattr(self, "_mb_off", None) is None:
            self._mb_off = (os.environ.get("HV_MESH_P2P", "1") == "0" or os.environ.get("HV_MESH_FLAGS", "1") == "0"
                            or self.local.device.type != "cuda" or dist.get_backend() != "nccl" or self.world_size > 16)
        if self._mb_off or getattr(self, "_p2p_off", False):
            return None
        if getattr(self, "_mb", None) is not None:
            return self._mb
        from . import marching_cubes as mc
        tail = 8  # >= McPlan.SLAB_TAIL, keeps the next array 32-byte aligned
        assert mc.McPlan.SLAB_TAIL <= tail
        a = lambda v: (v + 255) & ~255  # noqa: E731
        self._mb_halo = 0                                     # [parity 2][side 2][tsdf, colour][n_cols] float32
        self._mb_counts = a(self._mb_halo + 8 * n_cols * 4)   # [parity 2][3 * n_cols + tail] int32
        self._mb_count_words = 3 * n_cols + tail
        self._mb_flags = a(self._mb_counts + 2 * self._mb_count_words * 4)
        mb = peer.Mailboxes(dist, self.rank, self.world_size, self.local.device, self._mb_flags + 4 * self._FLAG_WORDS)
        if not mb.ok:
            self._mb_off = True
            return None
        dev, own = self.local.device, mb.base[self.rank]
        self._mb_err = torch.zeros(1, dtype=torch.int32, device=dev)
        self._mb_tails = torch.zeros((self.world_size, mc.McPlan.SLAB_TAIL), dtype=torch.int32, device=dev)
        self._mb_views = []
        for par in range(2):
            halo = {}
            for side, name in ((0, "lo"), (1, "hi")):
                if (side == 0 and self.rank == 0) or (side == 1 and self.rank == self.world_size - 1):
                    continue
                o = own + self._mb_halo + ((par * 2 + side) * 2) * n_cols * 4
                shape = tuple(int(d) for d in self.local._local_dim[:2])
                halo[f"tsdf_{name}"] = peer.view(o, shape, torch.float32, dev)
                halo[f"color_{name}"] = peer.view(o + n_cols * 4, shape, torch.float32, dev)
            counts = peer.view(own + self._mb_counts + par * self._mb_count_words * 4, (3 * n_cols + mc.McPlan.SLAB_TAIL,),
                               torch.int32, dev)
            self._mb_views.append((halo, counts))
        self._mb, self._epoch = mb, 0
        return mb

    def _get_mesh_peer(self, mark, trace, redo=False):
        """``get_mesh`` without a collective: seam planes are stored into the neighbours' mailboxes, every slab reads the
        others' per-column counts in place, the rows go straight into rank 0's HBM; flags (32-bit epochs stored into the
        receiver's memory behind the data, ``hv_peer_signal`` / ``hv_peer_wait``) order it all.  Two parities of every inbox:
        a slab can be at most one call ahead of its neighbours."""
        from . import marching_cubes as mc
        loc, torch, mb = self.local, self._torch, self._mb
        lib, rank, world = _lib_handle(), self.rank, self.world_size
        X, Y, Z = (int(d) for d in loc._local_dim)
        n_cols, z0, Zg = X * Y, self.ranges[self.rank][0], self.global_dim[2]
        self._epoch += 1
        e, par = self._epoch, self._epoch & 1
        halo, counts = self._mb_views[par]
        stream = torch.cuda.current_stream().cuda_stream
        flag = lambda w, i: mb.base[w] + self._mb_flags + 4 * i  # noqa: E731
        with torch.cuda.device(loc.device):
            # my first plane is the hi halo of the slab below, my last plane the lo halo of the slab above
            to_signal = []
            for nb, side, z in ((rank - 1, 1, 0), (rank + 1, 0, Z - 1)):
                if 0 <= nb < world:
                    dst = mb.base[nb] + self._mb_halo + ((par * 2 + side) * 2) * n_cols * 4
                    _check(lib.hv_copy_planes(_p(loc._tsdf_vol_gpu), _p(loc._color_vol_gpu), int(loc._pitch), z, n_cols,
                                              _c(dst), _c(dst + n_cols * 4), _c(stream)))
                    to_signal.append(flag(nb, side))
            mb.signal(to_signal, e, stream)
            first = 0 if rank > 0 else 1
            mb.wait(self._mb_flags + 4 * first, (1 if rank > 0 else 0) + (1 if rank < world - 1 else 0), e, self._mb_err, stream)
            mark("seam planes stored + arrived")
            plan = getattr(self, "_mc_plan_peer", None)
            if plan is None or plan.tsdf is not loc._tsdf_vol_gpu:
                plan = self._mc_plan_peer = mc.McPlan(loc._tsdf_vol_gpu, loc._color_vol_gpu, (X, Y, Z), loc._pitch, halo=halo,
                                                      z_global_offset=z0, z_global_dim=Zg)
            plan.halo = halo
            # the count is not waited for: its totals travel with the column counts, and the one host read below tells every
            # slab the sizes of all slabs -- and whether some slab's work records did not fit (then all go round again)
            plan.count(*getattr(self, "_mc_caps", (None, None)))
            plan._records_incomplete = False
            plan.column_counts(out=counts)
            mb.signal([flag(w, 8 + rank) for w in range(world)], e, stream)
            mb.wait(self._mb_flags + 4 * 8, world, e, self._mb_err, stream)
            mark("count + column counts ready everywhere")
            shifts = plan.column_shifts_peers([mb.base[w] + self._mb_counts + par * self._mb_count_words * 4 for w in range(world)],
                                              rank, self._mb_tails)
            tails = self._read_tails(torch.cat((self._mb_tails.view(-1), self._mb_err)))
            if int(tails[-1]):
                raise RuntimeError(f"sharded get_mesh: rank {rank} waited ~30 s for flag {int(tails[-1]) - 1} of its mailbox")
            tails = tails[:-1].reshape(world, plan.SLAB_TAIL)
            n_verts, n_tris = (int(v) for v in tails[rank, 2:4])
            self._mc_caps = (n_verts + n_verts // 2 + 1024, n_tris + n_tris // 2 + 1024)  # next time: records sized by this surface
            if tails[:, 5].any():
                assert not redo, "the work records did not fit twice in a row"
                return self._get_mesh_peer(mark, trace, redo=True)
            mark("shifts + tails read")
            vmin = min(mc._key_to_float(int(k)) for k in tails[:, 0])
            vmax = max(mc._key_to_float(int(k)) for k in tails[:, 1])
            n_verts_g, n_tris_g, n_edge_g = (int(v) for v in tails[:, 2:5].sum(axis=0))
            mc.check_surface(0.0, vmin, vmax, n_verts_g)
            buf = self._shared_mesh_buffer(n_verts_g, n_tris_g)
            assert buf is not None, "the mesh buffer could not be shared although the mailboxes could"
            offs, _ = peer.mesh_layout(n_verts_g, n_tris_g)
            plan.emit_direct(n_verts, n_tris, loc._voxel_size, loc._vol_origin, shifts, n_edge_g, n_verts_g, n_tris_g,
                             {name: buf.base + o for name, o in offs.items()})
            mb.signal([flag(0, 24 + rank)], e, stream)
            mark("emit into rank 0")
            self.last_mesh_trace, self.last_mesh_path = trace, "peer"
            if rank != 0:
                return None, None, None, None
            mb.wait(self._mb_flags + 4 * 24, world, e, self._mb_err, stream)
            mark("arrival")
            out = self._download_shared_mesh(buf, offs, n_verts_g, n_tris_g)
            if int(self._mb_err.item()):
                raise RuntimeError("sharded get_mesh: rank 0 waited ~30 s for the rows of a slab")
            mark("download")
            return out

    # -- the whole mesh in rank 0's HBM, written by every slab (hive_b200/peer.py)
    def _shared_mesh_buffer(self, n_verts_g, n_tris_g):
        """The buffer all slabs emit into, or None when the ranks cannot share memory (CPU process groups, HV_MESH_P2P=0,
        CUDA IPC unavailable): every rank takes the same decision -- the sizes come from the all-gathered totals and
        ``SharedBuffer`` agrees on success collectively.  Grown (collectively) when the mesh outgrows it."""
        dist, torch = self._dist, self._torch
        if getattr(self, "_p2p_off", None) is None:
            self._p2p_off = (os.environ.get("HV_MESH_P2P", "1") == "0" or self.local.device.type != "cuda"
                             or dist.get_backend() != "nccl")
        if self._p2p_off:
            return None
        _, need = peer.mesh_layout(n_verts_g, n_tris_g)
        buf = getattr(self, "_mesh_buf", None)
        if buf is not None and buf.nbytes >= need:
            return buf
        if buf is not None:
            buf.close()
        buf = peer.SharedBuffer(dist, self.rank, self.local.device, need + need // 2)
        if not buf.ok:
            self._p2p_off, self._mesh_buf = True, None
            return None
        self._mesh_buf = buf
        self._mesh_done = torch.zeros(1, dtype=torch.int32, device=self.local.device)
        return buf

    def _download_shared_mesh(self, buf, offs, n_verts, n_tris):
        torch = self._torch
        host = {"verts": torch.empty((n_verts, 3), dtype=torch.float32, pin_memory=True),
                "faces": torch.empty((n_tris, 3), dtype=torch.int32, pin_memory=True),
                "normals": torch.empty((n_verts, 3), dtype=torch.float32, pin_memory=True),
                "colors": torch.empty((n_verts, 3), dtype=torch.uint8, pin_memory=True)}
        with torch.cuda.device(self.local.device):
            stream = torch.cuda.current_stream()
            for name, h in host.items():
                if h.numel():
                    buf.copy_out(offs[name], h, stream.cuda_stream)
            stream.synchronize()
        return tuple(host[k].numpy() for k in ("verts", "faces", "normals", "colors"))

    def close(self):
        """Collective: releases the shared mesh buffer (otherwise it lives until the processes exit)."""
        buf = getattr(self, "_mesh_buf", None)
        if buf is not None:
            buf.close()
            self._mesh_buf = None
        mb = getattr(self, "_mb", None)
        if mb is not None:
            self._mb_views, self._mc_plan_peer = None, None  # (views into the mailbox)
            mb.close()
            self._mb = None


def _lib_handle():
    from . import _lib
    return _lib.load()


def _check(code):
    from . import _lib
    return _lib.check(code)


def _p(t):
    import ctypes
    return ctypes.c_void_p(t.data_ptr())


def _c(address):
    import ctypes
    return ctypes.c_void_p(int(address))


def exchange_halo_planes(dist, rank, world_size, planes, out=None):
    """Seam exchange: every rank sends its first plane down and its last plane up.  ``planes`` maps a name to
    (first, last); returns {name_lo, name_hi} tensors (absent at the ends of the stack), received into the tensors of
    ``out`` when given (an earlier result)."""
    import torch
    ops, out = [], (out if out is not None else {})
    for name, (first, last) in planes.items():
        if rank > 0:
            if f"{name}_lo" not in out:
                out[f"{name}_lo"] = torch.empty_like(first)
            ops.append(dist.P2POp(dist.isend, first, rank - 1))
            ops.append(dist.P2POp(dist.irecv, out[f"{name}_lo"], rank - 1))
        if rank < world_size - 1:
            if f"{name}_hi" not in out:
                out[f"{name}_hi"] = torch.empty_like(last)
            ops.append(dist.P2POp(dist.isend, last, rank + 1))
            ops.append(dist.P2POp(dist.irecv, out[f"{name}_hi"], rank + 1))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return out


_PART_FIELDS = (("verts", 3, "float32"), ("normals", 3, "float32"), ("colors", 3, "uint8"), ("vert_dst", 0, "int32"),
                ("faces", 3, "int32"), ("tri_dst", 0, "int32"))


def exchange_column_counts(dist, world_size, counts):
    """All-gather of ``McPlan.column_counts()`` (per-column counts + the slab's totals) -> (world_size, ...)."""
    import torch
    out = torch.empty((world_size,) + tuple(counts.shape), dtype=counts.dtype, device=counts.device)
    if counts.is_cuda:
        dist.all_gather_into_tensor(out.view(-1), counts.contiguous().view(-1))  # NCCL: one flat gather, no staging
    else:
        dist.all_gather(list(out.unbind(0)), counts.contiguous())  # gloo (CPU tests)
    return out


def _part_layout(nv, nf):
    """Byte offsets of the arrays of one slab mesh inside its packed message (every array starts on a 16-byte boundary)."""
    import torch
    offs, o = {}, 0
    for name, cols, dtype in _PART_FIELDS:
        n = nf if name in ("faces", "tri_dst") else nv
        nbytes = n * max(cols, 1) * torch.empty((), dtype=getattr(torch, dtype)).element_size()
        offs[name] = (o, nbytes, n, cols, dtype)
        o = (o + nbytes + 15) & ~15
    return offs, o


def gather_placed_parts(dist, rank, world_size, part, device, all_sizes):
    """Variable-length gather of the per-slab meshes to rank 0 -- ``all_sizes[w]`` = (vertices, triangles) of slab w, known
    to everyone from the column-count exchange.  Every slab packs its six arrays into ONE message; rank 0 posts all
    receives at once and unpacks by view (nothing waits for anything else, nothing is copied twice)."""
    import torch
    if rank != 0:
        offs, total = _part_layout(int(part["vert_dst"].shape[0]), int(part["tri_dst"].shape[0]))
        if total:
            msg = torch.empty(total, dtype=torch.uint8, device=device)
            for name, (o, nbytes, _, _, _) in offs.items():
                if nbytes:
                    msg[o:o + nbytes] = part[name].contiguous().view(-1).view(torch.uint8)
            for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, msg, 0)]):
                req.wait()
        return None
    ops, msgs = [], []
    for src in range(1, world_size):
        nv, nf = (int(v) for v in all_sizes[src])
        offs, total = _part_layout(nv, nf)
        msg = torch.empty(total, dtype=torch.uint8, device=device)
        if total:
            ops.append(dist.P2POp(dist.irecv, msg, src))
        msgs.append((offs, msg))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    parts = [part]
    for offs, msg in msgs:
        got = {}
        for name, (o, nbytes, n, cols, dtype) in offs.items():
            t = msg[o:o + nbytes].view(getattr(torch, dtype))
            got[name] = t.view(n, cols) if cols else t
        parts.append(got)
    return parts


def place_mesh_parts(parts, n_verts, n_tris, scatter_rows):
    """Rank 0: every row of every slab mesh goes where its slab said it goes (``vert_dst`` / ``tri_dst``).
    ``scatter_rows(src, index, dst)`` is ``hive_b200.marching_cubes.scatter_rows`` (the CPU tests pass a numpy stand-in)."""
    import torch
    dev = parts[0]["verts"].device
    verts = torch.empty((n_verts, 3), dtype=torch.float32, device=dev)
    normals = torch.empty((n_verts, 3), dtype=torch.float32, device=dev)
    colors = torch.empty((n_verts, 3), dtype=torch.uint8, device=dev)
    faces = torch.empty((n_tris, 3), dtype=torch.int32, device=dev)
    assert sum(int(p["vert_dst"].shape[0]) for p in parts) == n_verts, "the slabs' vertices do not add up to the whole mesh"
    assert sum(int(p["tri_dst"].shape[0]) for p in parts) == n_tris, "the slabs' triangles do not add up to the whole mesh"
    for p in parts:
        scatter_rows(p["verts"], p["vert_dst"], verts)
        scatter_rows(p["normals"], p["vert_dst"], normals)
        scatter_rows(p["colors"], p["vert_dst"], colors)
        scatter_rows(p["faces"], p["tri_dst"], faces)
    return verts, faces, normals, colors
