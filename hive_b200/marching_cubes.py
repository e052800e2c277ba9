"""Marching-cubes mesh extraction on the GPU: host side of ``hv_mc_count`` / ``hv_mc_emit``.

Replaces ``skimage.measure.marching_cubes(tsdf_vol, level=0)`` plus the post-processing of
``TSDFVolume.get_mesh`` (third_party/tsdf_fusion_python/fusion.py:332-350).  ``marching_cubes`` mirrors the skimage
signature HIVE relies on (volume, level -> verts, faces, normals, values; ``ValueError`` when the level is outside the
data range -- ``scripts/experiments.py:167`` catches exactly that -- and ``RuntimeError`` when no surface is found).
There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
import threading
from dataclasses import dataclass

import numpy as np

from . import _lib
from . import _xfer


#: kernels launched by one extraction (sign, detect, count, scan, emit); bench.py reports it
LAUNCHES_PER_EXTRACTION = 5


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _key_to_float(key: int) -> float:
    """Inverse of the order-preserving int32 key the kernels use for the min/max reduction."""
    key = int(key) & 0xFFFFFFFF
    if key >= 0x80000000:
        key -= 1 << 32
    bits = key if key >= 0 else key ^ 0x7FFFFFFF
    return float(np.array([bits & 0xFFFFFFFF], dtype=np.uint32).view(np.float32)[0])


@dataclass
class DeviceMesh:
    """A mesh resident in HBM (torch tensors)."""
    verts: object  # (V, 3) float32 world coordinates
    verts_voxel: object  # (V, 3) float32 voxel units (what skimage returns), or None
    normals: object  # (V, 3) float32
    colors: object  # (V, 3) uint8 r, g, b
    faces: object  # (T, 3) int32 indices, or None when face_keys is used
    face_keys: object  # (T, 3) int64 global vertex keys, or None
    face_cubes: object  # (T,) int64 global cube index of each face (with face_keys), or None
    vert_keys: object  # (V,) int64 global vertex keys, ascending
    vmin: float
    vmax: float

    def to_host(self):
        return _xfer.to_host(self.verts, self.faces, self.normals, self.colors)


_PINNED = {}


def _pinned_totals(torch, device):
    key = (threading.get_ident(), str(device))
    buf = _PINNED.get(key)
    if buf is None:
        buf = _PINNED[key] = torch.empty(8, dtype=torch.int64).pin_memory()
    return buf


class McPlan:
    """Scratch + launch state of one extraction over planes ``[X][Y][pitch]`` (optionally a z-slab with halo planes)."""

    def __init__(self, tsdf, color, dims, pitch, halo=None, z_global_offset=0, z_global_dim=None, level=0.0):
        torch = _lib.require_cuda()
        self._torch = torch
        self._lib = _lib.load()
        self.tsdf, self.color = tsdf, color
        self.dims = (ctypes.c_int64 * 3)(*[int(d) for d in dims])
        self.pitch = int(pitch)
        self.level = float(level)
        self.halo = halo or {}  # keys: tsdf_lo, tsdf_hi, color_lo, color_hi -> (X, Y) float32 device tensors
        self.z_global_offset = int(z_global_offset)
        self.z_global_dim = int(dims[2]) if z_global_dim is None else int(z_global_dim)
        nbytes = ctypes.c_int64(0)
        _lib.check(self._lib.hv_mc_scratch_bytes(ctypes.cast(self.dims, ctypes.c_void_p),
                                                 int(self.halo.get("tsdf_lo") is not None),
                                                 int(self.halo.get("tsdf_hi") is not None), ctypes.byref(nbytes)))
        self.scratch_bytes = int(nbytes.value)
        self.device = tsdf.device
        with torch.cuda.device(self.device):
            self.scratch = torch.empty(self.scratch_bytes + 256, dtype=torch.uint8, device=self.device)
            pad = (-self.scratch.data_ptr()) % 256
            self.scratch_ptr = ctypes.c_void_p(self.scratch.data_ptr() + pad)
            self.totals = torch.zeros(8, dtype=torch.int64, device=self.device)
            _lib.check(self._lib.hv_mc_scratch_init(ctypes.cast(self.dims, ctypes.c_void_p),
                                                    int(self.halo.get("tsdf_lo") is not None),
                                                    int(self.halo.get("tsdf_hi") is not None), self.scratch_ptr,
                                                    self.scratch_bytes,
                                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
        self.work = None
        self.work_ptr = None
        self.rec_verts = self.rec_cubes = 0
        self._records_incomplete = False

    def _ensure_work(self, rec_verts, rec_cubes):
        """Work records of the emit kernels (one per edge vertex, one per surface cube): a surface cube has at least one
        triangle, so (vertex capacity, triangle capacity) always suffices."""
        rec_verts, rec_cubes = int(rec_verts), int(rec_cubes)
        if self.work is not None and rec_verts <= self.rec_verts and rec_cubes <= self.rec_cubes:
            return
        torch = self._torch
        nbytes = ctypes.c_int64(0)
        _lib.check(self._lib.hv_mc_work_bytes(rec_verts, rec_cubes, ctypes.byref(nbytes)))
        with torch.cuda.device(self.device):
            self.work = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=self.device)
        self.work_ptr = ctypes.c_void_p(self.work.data_ptr() + (-self.work.data_ptr()) % 256)
        self.rec_verts, self.rec_cubes = rec_verts, rec_cubes

    def default_capacity(self):
        """A few voxel layers' worth of surface (an extraction that needs more counts twice)."""
        X, Y, Z = (int(d) for d in self.dims)
        area = X * Y + Y * Z + X * Z
        return min(6 * area, 3 * X * Y * Z), min(12 * area, 5 * X * Y * Z)

    def count(self, cap_verts=None, cap_tris=None):
        """Launch the classification (sign bits, per-group counts, work records for a surface of at most ``cap_verts``
        vertices / ``cap_tris`` triangles); returns the device tensor of totals (no sync)."""
        torch = self._torch
        if cap_verts is None or cap_tris is None:
            cap_verts, cap_tris = self.default_capacity()
        self._ensure_work(cap_verts, cap_tris)
        with torch.cuda.device(self.device):
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self._lib.hv_mc_count(
                _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")),
                ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                self.work_ptr, self.rec_verts, self.rec_cubes, _ptr(self.totals), stream))
        return self.totals

    def read_totals(self):
        """The one host round trip of an extraction (output sizes are data dependent): async D2H of the four counters
        into a cached pinned buffer, then a stream synchronise (a pageable ``.cpu()`` costs several times more)."""
        torch = self._torch
        host = _pinned_totals(torch, self.device)
        with torch.cuda.device(self.device):
            host.copy_(self.totals, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        vals = [int(v) for v in host.tolist()]
        n_verts, n_tris, kmin, kmax = vals[:4]
        self._records_incomplete = bool(vals[6])
        return n_verts, n_tris, _key_to_float(kmin), _key_to_float(kmax)

    def emit(self, n_verts, n_tris, voxel_size, origin, want_voxel=False, keys=False) -> DeviceMesh:
        torch = self._torch
        dev = self.device
        origin = np.ascontiguousarray(origin, dtype=np.float32)
        if self._records_incomplete:  # the surface outgrew the work buffer of the count: count again, exactly sized
            self.count(n_verts, n_tris)
            self._records_incomplete = False
        with torch.cuda.device(dev):
            verts = torch.empty((n_verts, 3), dtype=torch.float32, device=dev)
            verts_voxel = torch.empty((n_verts, 3), dtype=torch.float32, device=dev) if want_voxel else None
            normals = torch.empty((n_verts, 3), dtype=torch.float32, device=dev)
            colors = torch.empty((n_verts, 3), dtype=torch.uint8, device=dev)
            vert_keys = torch.empty((max(n_verts, 1),), dtype=torch.int64, device=dev) if keys else None
            faces = None if keys else torch.empty((n_tris, 3), dtype=torch.int32, device=dev)
            face_keys = torch.empty((n_tris, 3), dtype=torch.int64, device=dev) if keys else None
            face_cubes = torch.empty((n_tris,), dtype=torch.int64, device=dev) if keys else None
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self._lib.hv_mc_emit(
                _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")), _ptr(self.color),
                _ptr(self.halo.get("color_lo")), _ptr(self.halo.get("color_hi")),
                ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                self.work_ptr, self.rec_verts, self.rec_cubes,
                self.z_global_offset, self.z_global_dim, ctypes.c_float(np.float32(voxel_size)),
                origin.ctypes.data_as(ctypes.c_void_p), n_verts, n_tris, _ptr(verts), _ptr(verts_voxel), _ptr(normals),
                _ptr(colors), _ptr(faces), _ptr(face_keys), _ptr(face_cubes), _ptr(vert_keys), _ptr(self.totals), stream))
        return DeviceMesh(verts, verts_voxel, normals, colors, faces, face_keys, face_cubes,
                          vert_keys[:n_verts] if keys else None, 0.0, 0.0)


    # ------------------------------------------------------------------ z-slab placement (hive_b200/sharded.py)
    def ensure_records(self, n_verts, n_tris):
        """Make the work records of the last count complete (count again, exactly sized, if they did not fit)."""
        if self._records_incomplete:
            self.count(n_verts, n_tris)
            self._records_incomplete = False

    #: words appended to a slab's column counts: min key, max key, vertices, triangles, edge vertices, 1 if the work records
    #: of the count did not fit (the counts are then not valid: count again with the sizes given here)
    SLAB_TAIL = 6

    def column_counts(self, out=None):
        """int32 (3 * X*Y + SLAB_TAIL,) device tensor: edge vertices / triangles / centre vertices this slab owns per
        column, then the slab's totals (everything the other slabs and rank 0 need to know about it -- one all-gather, or
        read in place by the other slabs when ``out`` lies in memory they have mapped)."""
        torch = self._torch
        X, Y = int(self.dims[0]), int(self.dims[1])
        with torch.cuda.device(self.device):
            counts = out if out is not None else torch.empty((3 * X * Y + self.SLAB_TAIL,), dtype=torch.int32, device=self.device)
            counts[3 * X * Y:] = self.totals.view(torch.int32)[self._tail_words(torch)]  # low words of totals[2, 3, 0, 1, 4, 6]
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self._lib.hv_mc_column_counts(
                _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")),
                ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                self.work_ptr, self.rec_verts, self.rec_cubes, _ptr(self.totals), _ptr(counts), stream))
        return counts

    def _tail_words(self, torch):
        idx = getattr(self, "_tail_idx", None)
        if idx is None:
            idx = self._tail_idx = torch.tensor([4, 6, 0, 2, 8, 12], dtype=torch.int64, device=self.device)
        return idx

    def column_shifts(self, all_counts, rank):
        """``all_counts``: int32 (world, 3 * X*Y + SLAB_TAIL), ``column_counts()`` of every slab.  Returns (shifts int32
        (4, X*Y), totals of the whole volume int64 (4,): edge vertices, triangles, centre vertices) -- device tensors, no
        sync."""
        torch = self._torch
        world, stride = (int(v) for v in all_counts.shape)
        n_cols = (stride - self.SLAB_TAIL) // 3
        nbytes = ctypes.c_int64(0)
        _lib.check(self._lib.hv_mc_shift_scratch_bytes(n_cols, ctypes.byref(nbytes)))
        with torch.cuda.device(self.device):
            scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=self.device)
            sptr = ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256)
            shifts = torch.empty((4, n_cols), dtype=torch.int32, device=self.device)
            g_totals = torch.zeros(4, dtype=torch.int64, device=self.device)
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self._lib.hv_mc_column_shifts(_ptr(all_counts.contiguous()), stride, world, int(rank), n_cols, _ptr(shifts),
                                                     _ptr(g_totals), sptr, int(nbytes.value), stream))
            scratch.record_stream(torch.cuda.current_stream())
        return shifts, g_totals

    def column_shifts_peers(self, slab_addresses, rank, tails):
        """``column_shifts`` reading every slab's ``column_counts(out=...)`` where it is: ``slab_addresses[w]`` is this
        process's address of slab w's counts (mapped peer memory).  ``tails`` (int32 device tensor (world, SLAB_TAIL))
        receives the slabs' tail words.  Returns the shifts; no sync."""
        torch = self._torch
        world = len(slab_addresses)
        n_cols = int(self.dims[0]) * int(self.dims[1])
        nbytes = ctypes.c_int64(0)
        _lib.check(self._lib.hv_mc_shift_scratch_bytes(n_cols, ctypes.byref(nbytes)))
        with torch.cuda.device(self.device):
            scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=self.device)
            sptr = ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256)
            shifts = torch.empty((4, n_cols), dtype=torch.int32, device=self.device)
            g_totals = torch.zeros(4, dtype=torch.int64, device=self.device)
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            ptrs = (ctypes.c_void_p * world)(*[ctypes.c_void_p(int(a)) for a in slab_addresses])
            _lib.check(self._lib.hv_mc_column_shifts_peers(ptrs, world, int(rank), n_cols, self.SLAB_TAIL, _ptr(shifts),
                                                           _ptr(g_totals), _ptr(tails), sptr, int(nbytes.value), stream))
            scratch.record_stream(torch.cuda.current_stream())
        return shifts

    def emit_placed(self, n_verts, n_tris, voxel_size, origin, shifts, n_edge_global):
        """The slab's mesh with faces as indices into the mesh of the WHOLE volume; returns a dict of device tensors:
        verts, normals, colors, faces, vert_dst (where each vertex goes in the whole mesh), tri_dst."""
        torch = self._torch
        dev = self.device
        origin = np.ascontiguousarray(origin, dtype=np.float32)
        self.ensure_records(n_verts, n_tris)
        with torch.cuda.device(dev):
            out = {"verts": torch.empty((n_verts, 3), dtype=torch.float32, device=dev),
                   "normals": torch.empty((n_verts, 3), dtype=torch.float32, device=dev),
                   "colors": torch.empty((n_verts, 3), dtype=torch.uint8, device=dev),
                   "faces": torch.empty((n_tris, 3), dtype=torch.int32, device=dev),
                   "vert_dst": torch.empty((n_verts,), dtype=torch.int32, device=dev),
                   "tri_dst": torch.empty((n_tris,), dtype=torch.int32, device=dev)}
            if n_verts == 0 and n_tris == 0:
                return out
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self._lib.hv_mc_emit_placed(
                _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")), _ptr(self.color),
                _ptr(self.halo.get("color_lo")), _ptr(self.halo.get("color_hi")),
                ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                self.work_ptr, self.rec_verts, self.rec_cubes, self.z_global_offset, self.z_global_dim,
                ctypes.c_float(np.float32(voxel_size)), origin.ctypes.data_as(ctypes.c_void_p), n_verts, n_tris,
                _ptr(out["verts"]), _ptr(out["normals"]), _ptr(out["colors"]), _ptr(out["faces"]), _ptr(shifts),
                int(n_edge_global), _ptr(out["vert_dst"]), _ptr(out["tri_dst"]), _ptr(self.totals), stream))
        return out

    def emit_direct(self, n_verts, n_tris, voxel_size, origin, shifts, n_edge_global, whole_verts, whole_tris, addr):
        """The slab's rows written at their positions in the mesh of the WHOLE volume (``hv_mc_emit_direct``): ``addr`` maps
        verts / normals / colors / faces to the device addresses (ints) of the whole mesh's arrays -- possibly another
        GPU's memory mapped into this process (``hive_b200.peer``), in which case the stores travel over NVLink."""
        torch = self._torch
        origin = np.ascontiguousarray(origin, dtype=np.float32)
        self.ensure_records(n_verts, n_tris)
        if n_verts == 0 and n_tris == 0:
            return
        with torch.cuda.device(self.device):
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self._lib.hv_mc_emit_direct(
                _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")), _ptr(self.color),
                _ptr(self.halo.get("color_lo")), _ptr(self.halo.get("color_hi")),
                ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                self.work_ptr, self.rec_verts, self.rec_cubes, self.z_global_offset, self.z_global_dim,
                ctypes.c_float(np.float32(voxel_size)), origin.ctypes.data_as(ctypes.c_void_p), n_verts, n_tris,
                _ptr(shifts), int(n_edge_global), int(whole_verts), int(whole_tris), ctypes.c_void_p(addr["verts"]),
                ctypes.c_void_p(addr["normals"]), ctypes.c_void_p(addr["colors"]), ctypes.c_void_p(addr["faces"]),
                _ptr(self.totals), stream))

    def extract(self, voxel_size, origin, cap_verts, cap_tris, want_voxel=False) -> DeviceMesh:
        """count + emit in one library call (``hv_mc_extract``) into buffers of the given capacity; falls back to an
        exactly sized ``emit`` when the surface turns out larger.  Whole volumes only (faces as indices)."""
        torch = self._torch
        dev = self.device
        origin = np.ascontiguousarray(origin, dtype=np.float32)
        host = _pinned_totals(torch, dev)
        self._ensure_work(cap_verts, cap_tris)
        with torch.cuda.device(dev):
            verts = torch.empty((cap_verts, 3), dtype=torch.float32, device=dev)
            verts_voxel = torch.empty((cap_verts, 3), dtype=torch.float32, device=dev) if want_voxel else None
            normals = torch.empty((cap_verts, 3), dtype=torch.float32, device=dev)
            colors = torch.empty((cap_verts, 3), dtype=torch.uint8, device=dev)
            faces = torch.empty((cap_tris, 3), dtype=torch.int32, device=dev)
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            rc = self._lib.hv_mc_extract(
                _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")), _ptr(self.color),
                _ptr(self.halo.get("color_lo")), _ptr(self.halo.get("color_hi")),
                ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                self.work_ptr, self.rec_verts, self.rec_cubes, self.z_global_offset, self.z_global_dim, ctypes.c_float(np.float32(voxel_size)),
                origin.ctypes.data_as(ctypes.c_void_p), cap_verts, cap_tris, _ptr(verts), _ptr(verts_voxel),
                _ptr(normals), _ptr(colors), _ptr(faces), None, None, None, _ptr(self.totals),
                ctypes.c_void_p(host.data_ptr()), stream)
        vals = [int(v) for v in host.tolist()]
        n_verts, n_tris, kmin, kmax = vals[:4]
        self._records_incomplete = bool(vals[6])
        vmin, vmax = _key_to_float(kmin), _key_to_float(kmax)
        if rc == _lib.HV_E_CAPACITY:
            mesh = self.emit(n_verts, n_tris, voxel_size, origin, want_voxel=want_voxel)
        else:
            _lib.check(rc)
            mesh = DeviceMesh(verts[:n_verts], verts_voxel[:n_verts] if want_voxel else None, normals[:n_verts],
                              colors[:n_verts], faces[:n_tris], None, None, None, 0.0, 0.0)
        mesh.vmin, mesh.vmax = vmin, vmax
        return mesh

    #: set by extract_overlapped(timing_event=True) when the emit was launched behind the count without a host round trip
    done_event = None

    def extract_overlapped(self, voxel_size, origin, cap_verts, cap_tris, want_voxel=False, speculative=False,
                           timing_event=False) -> DeviceMesh:
        """count -> (allocate the outputs while the classification runs) -> emit.  ``speculative``: the capacity comes from
        an earlier extraction of the same volume, so emit is launched right behind the classification and the totals are
        read afterwards -- no host round trip between the kernels; a surface that outgrew the capacity is emitted again
        into exactly sized buffers."""
        torch = self._torch
        dev = self.device
        self.count(cap_verts, cap_tris)
        with torch.cuda.device(dev):
            verts = torch.empty((cap_verts, 3), dtype=torch.float32, device=dev)
            verts_voxel = torch.empty((cap_verts, 3), dtype=torch.float32, device=dev) if want_voxel else None
            normals = torch.empty((cap_verts, 3), dtype=torch.float32, device=dev)
            colors = torch.empty((cap_verts, 3), dtype=torch.uint8, device=dev)
            faces = torch.empty((cap_tris, 3), dtype=torch.int32, device=dev)
        origin = np.ascontiguousarray(origin, dtype=np.float32)

        def launch():
            with torch.cuda.device(dev):
                stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
                _lib.check(self._lib.hv_mc_emit(
                    _ptr(self.tsdf), _ptr(self.halo.get("tsdf_lo")), _ptr(self.halo.get("tsdf_hi")), _ptr(self.color),
                    _ptr(self.halo.get("color_lo")), _ptr(self.halo.get("color_hi")),
                    ctypes.cast(self.dims, ctypes.c_void_p), self.pitch, self.level, self.scratch_ptr, self.scratch_bytes,
                    self.work_ptr, self.rec_verts, self.rec_cubes,
                    self.z_global_offset, self.z_global_dim, ctypes.c_float(np.float32(voxel_size)),
                    origin.ctypes.data_as(ctypes.c_void_p), cap_verts, cap_tris, _ptr(verts), _ptr(verts_voxel), _ptr(normals),
                    _ptr(colors), _ptr(faces), None, None, None, _ptr(self.totals), stream))

        done = None
        if speculative:
            launch()
            if timing_event:  # end of the device work of this extraction (what follows is the host reading the totals)
                with torch.cuda.device(dev):
                    done = torch.cuda.Event(enable_timing=True)
                    done.record()
        n_verts, n_tris, vmin, vmax = self.read_totals()
        check_surface(self.level, vmin, vmax, n_verts)
        if n_verts > cap_verts or n_tris > cap_tris or self._records_incomplete:
            mesh = self.emit(n_verts, n_tris, voxel_size, origin, want_voxel=want_voxel)
        else:
            if not speculative:
                launch()
            mesh = DeviceMesh(verts[:n_verts], verts_voxel[:n_verts] if want_voxel else None, normals[:n_verts],
                              colors[:n_verts], faces[:n_tris], None, None, None, 0.0, 0.0)
        mesh.vmin, mesh.vmax = vmin, vmax
        self.done_event = done
        return mesh


def scatter_rows(src, index, dst):
    """``dst[index[i]] = src[i]`` on the device (rows of 3 bytes or a multiple of 4): assembles a whole-volume mesh from slab
    meshes whose rows know where they go (``McPlan.emit_placed``)."""
    torch = _lib.require_cuda()
    n = int(src.shape[0])
    if n == 0:
        return
    row_bytes = src.element_size() * (int(src.numel()) // n)
    src = src.contiguous()
    assert dst.is_contiguous() and index.dtype == torch.int32 and index.is_contiguous()
    with torch.cuda.device(src.device):
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(_lib.load().hv_scatter_rows(_ptr(src), _ptr(index), n, row_bytes, _ptr(dst), stream))


def check_surface(level, vmin, vmax, n_verts):
    """The two errors of skimage.measure.marching_cubes that HIVE can observe."""
    if level < vmin or level > vmax:
        raise ValueError("Surface level must be within volume data range.")
    if n_verts == 0:
        raise RuntimeError("No surface found at the given iso value.")


#: surface sizes seen per volume shape: the capacity guess of the next extraction (1.5x the last result)
_LAST_SIZES = {}


def extract_device_mesh(vol, level=0.0, want_voxel=False, timing_event=False) -> DeviceMesh:
    """Mesh of a (single-GPU, whole) ``TSDFVolume`` as device tensors.  ``timing_event``: leave a CUDA event recorded
    right behind the last kernel in ``vol._mc_plan.done_event`` (bench.py times the device work with it)."""
    X, Y, Z = (int(d) for d in vol._local_dim)
    assert min(X, Y, Z) >= 2, "Input array must be at least 2x2x2."
    # scratch (1.1 B/voxel) is kept with the volume: repeated extractions (HIVE meshes once, a live preview meshes often)
    # skip the allocation and the classification can start immediately
    plan = getattr(vol, "_mc_plan", None)
    if plan is None or plan.level != float(level) or plan.tsdf is not vol._tsdf_vol_gpu:
        plan = vol._mc_plan = McPlan(vol._tsdf_vol_gpu, vol._color_vol_gpu, (X, Y, Z), vol._pitch, level=level)
    # capacity: a few voxel layers' worth of surface, or what the previous extraction of this shape needed
    guess = _LAST_SIZES.get((X, Y, Z))
    if guess is None:
        area = X * Y + Y * Z + X * Z
        guess = (6 * area, 12 * area)
    cap_v = min(guess[0], 3 * X * Y * Z)
    cap_t = min(guess[1], 5 * X * Y * Z)
    mesh = plan.extract_overlapped(vol._voxel_size, vol._vol_origin, cap_v, cap_t, want_voxel=want_voxel,
                                   speculative=(X, Y, Z) in _LAST_SIZES, timing_event=timing_event)
    n_v, n_t = int(mesh.verts.shape[0]), int(mesh.faces.shape[0])
    _LAST_SIZES[(X, Y, Z)] = (n_v + n_v // 2 + 1024, n_t + n_t // 2 + 1024)
    return mesh


def extract_mesh(vol, srgb_colors=False):
    """``TSDFVolume.get_mesh`` (fusion.py:332-350): verts (V,3) f32 world, faces (F,3) i32, norms (V,3) f32,
    colors (V,3) u8.  The marching cubes inside is ``marching_cubes`` of this module: parity with skimage unpinned
    (see there); everything after that call (fusion.py:339-348) is pinned against the reference's own lines.  ``srgb_colors`` applies the vertex-colour conversion of hive/pipeline.py:281-282 on the device,
    before the download."""
    mesh = extract_device_mesh(vol)
    if srgb_colors and mesh.colors.numel():
        from . import image_processing
        image_processing.apply_byte_lut_device(mesh.colors, image_processing.srgb_vertex_color_lut())
    return mesh.to_host()


def marching_cubes(volume, level=None, *, spacing=(1., 1., 1.), gradient_direction='descent', step_size=1,
                   allow_degenerate=True, method='lewiner', mask=None):
    """GPU drop-in for the ``skimage.measure.marching_cubes`` call of fusion.py:338 on a host array.

    Only the argument values the reference uses are supported (defaults, ``level`` given); returns
    ``(verts, faces, normals, values)`` with verts in voxel units, axis order (axis0, axis1, axis2).

    PARITY WITH SKIMAGE UNPINNED: scikit-image is installed on neither box and its tables cannot be consulted.  The
    surface is Lewiner's (face test, interior test, tunnel tilings: DESIGN.md 2.1), checked against an independently
    written restatement and the triangle counts per sub-case of SURVEY.md App. C; known differences from the real
    implementation: the triangulation inside a cube and the vertex numbering are this package's (vertices sorted by
    lattice edge, then the cube-centre vertices), and ``values`` is the constant ``level`` (skimage returns a local
    maximum of the data near each vertex; HIVE discards it, fusion.py:338).
    """
    torch = _lib.require_cuda()
    if (tuple(spacing) != (1., 1., 1.) or gradient_direction != 'descent' or step_size != 1 or not allow_degenerate
            or mask is not None or method != 'lewiner'):
        raise NotImplementedError("hive_b200.marching_cubes supports the reference's call only (fusion.py:338)")
    volume = np.asarray(volume)
    if volume.ndim != 3:
        raise ValueError("Input volume should be a 3D numpy array.")
    if min(volume.shape) < 2:
        raise ValueError("Input array must be at least 2x2x2.")
    vol32 = np.ascontiguousarray(volume, dtype=np.float32)
    if level is None:
        level = 0.5 * (float(vol32.min()) + float(vol32.max()))
    dev = torch.device("cuda", torch.cuda.current_device())
    tsdf = _xfer.to_device(vol32, dev)
    color = torch.zeros_like(tsdf)
    X, Y, Z = vol32.shape
    plan = McPlan(tsdf, color, (X, Y, Z), Z, level=float(level))
    plan.count()
    n_verts, n_tris, vmin, vmax = plan.read_totals()
    check_surface(float(level), vmin, vmax, n_verts)
    mesh = plan.emit(n_verts, n_tris, 1.0, np.zeros(3, np.float32), want_voxel=True)
    values = np.full(n_verts, np.float32(level))
    verts, faces, normals = _xfer.to_host(mesh.verts_voxel, mesh.faces, mesh.normals)
    return verts, faces, normals, values
