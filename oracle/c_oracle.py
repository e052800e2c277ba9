"""ORACLE -- test infrastructure, not product code.

ctypes front-end of ``oracle/libhive_oracle.so`` (the plain-C restatements in this directory).  Only ``tests/``,
``__graft_entry__`` and ``bench.py``'s CPU-baseline legs may import it.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from . import fusion_oracle as fo

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    """Compile the C restatements (gcc, a second or two)."""
    so = os.path.join(_HERE, "libhive_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("tsdf_oracle.c", "mc_oracle.c")]
    stale = force or not os.path.exists(so) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        subprocess.run(["make", "-s", "-B", "-C", _HERE, "libhive_oracle.so"], check=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.hvo_integrate.restype = ctypes.c_longlong
        _LIB.hvo_num_threads.restype = ctypes.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def num_threads() -> int:
    return int(lib().hvo_num_threads())


def use_all_host_threads() -> int:
    """OpenMP threads = all host cores, whatever OMP_NUM_THREADS says (torchrun sets it to 1 for every rank)."""
    lib().hvo_set_num_threads(ctypes.c_int(os.cpu_count() or 1))
    return num_threads()


class COracleTSDFVolume:
    """Same contract as ``fusion_oracle.OracleTSDFVolume`` (fusion.py:21-303), executed by tsdf_oracle.c."""

    def __init__(self, vol_bnds, voxel_size):
        self._vol_bnds, self._vol_dim, self._vol_origin = fo.volume_geometry(vol_bnds, voxel_size)
        self._voxel_size = float(voxel_size)
        self._trunc_margin = 5 * self._voxel_size
        self._tsdf_vol_cpu = np.ones(self._vol_dim, dtype=np.float32)
        self._weight_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self._color_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self._tables = fo.world_tables(self._vol_origin, self._voxel_size, self._vol_dim)
        self.last_updated = 0

    def integrate(self, color_im, depth_im, cam_intr, cam_pose, obs_weight=1.0):
        self.integrate_prepared(prepare_frame(color_im, depth_im, cam_intr, cam_pose), obs_weight)

    def integrate_prepared(self, frame, obs_weight=1.0):
        """``frame`` = prepare_frame(...): lets a caller that integrates the same frames into many volumes (the
        block-sampled parity checks) pay the colour fold and the pose inverse once."""
        folded, depth, w2c, intr, H, W = frame
        X, Y, Z = (int(d) for d in self._vol_dim)
        wx, wy, wz = self._tables
        self.last_updated = int(lib().hvo_integrate(
            _p(self._tsdf_vol_cpu), _p(self._weight_vol_cpu), _p(self._color_vol_cpu),
            ctypes.c_longlong(X), ctypes.c_longlong(Y), ctypes.c_longlong(Z),
            _p(wx), _p(wy), _p(wz), _p(w2c), _p(intr), _p(depth), _p(folded),
            ctypes.c_int(H), ctypes.c_int(W),
            ctypes.c_double(self._trunc_margin), ctypes.c_double(float(obs_weight))))

    def get_volume(self):
        return self._tsdf_vol_cpu, self._color_vol_cpu


def prepare_frame(color_im, depth_im, cam_intr, cam_pose):
    """Per-frame inputs of hvo_integrate: folded colour (fusion.py:231-232), depth as given (the CPU path keeps
    depth_im's precision), inv(cam_pose)[:3] (fusion.py:263), float32 intrinsics (fusion.py:196)."""
    H, W = depth_im.shape
    folded = np.ascontiguousarray(fo.fold_color(color_im), dtype=np.float32)
    depth = np.ascontiguousarray(depth_im, dtype=np.float64)
    w2c = np.ascontiguousarray(np.linalg.inv(np.asarray(cam_pose, dtype=np.float64))[:3, :])
    k = np.asarray(cam_intr).astype(np.float32)
    intr = np.array([k[0, 0], k[1, 1], k[0, 2], k[1, 2]], dtype=np.float32)
    return folded, depth, w2c, intr, H, W


class COracleBlock(COracleTSDFVolume):
    """A box of a larger volume: the caller passes the box's slices of the GLOBAL world-coordinate tables
    (float32(float64(origin) + voxel_size * global_index), fusion.py:183-188), so the oracle sees exactly the voxels
    the product holds at those indices -- how volumes too large for the host (2048^3) are checked block by block."""

    def __init__(self, tables, voxel_size):
        self._tables = tuple(np.ascontiguousarray(t, dtype=np.float32) for t in tables)
        self._vol_dim = np.array([len(t) for t in self._tables])
        self._voxel_size = float(voxel_size)
        self._trunc_margin = 5 * self._voxel_size
        self._tsdf_vol_cpu = np.ones(self._vol_dim, dtype=np.float32)
        self._weight_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self._color_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self.last_updated = 0


def marching_cubes(volume, level=0.0, with_keys=False, with_cubes=False):
    """``skimage.measure.marching_cubes(volume, level)`` as restated by mc_oracle.c (fusion.py:338; PARITY UNPINNED,
    see the header of that file).  Returns (verts f32 (V,3) in voxel units, faces i32 (F,3), normals f32 (V,3),
    values f32 (V,)) like skimage; ``with_keys`` appends the canonical edge key of every vertex."""
    vol = np.ascontiguousarray(volume, dtype=np.float32)
    if vol.ndim != 3:
        raise ValueError("Input volume should be a 3D numpy array.")
    if min(vol.shape) < 2:
        raise ValueError("Input array must be at least 2x2x2.")
    L = lib()
    L.hvo_marching_cubes.restype = ctypes.c_int
    L.hvo_mesh_nv.restype = ctypes.c_int64
    L.hvo_mesh_nf.restype = ctypes.c_int64
    handle = ctypes.c_void_p()
    rc = L.hvo_marching_cubes(_p(vol), ctypes.c_int64(vol.shape[0]), ctypes.c_int64(vol.shape[1]),
                              ctypes.c_int64(vol.shape[2]), ctypes.c_double(float(level)), ctypes.byref(handle))
    if rc == -1:
        raise ValueError("Surface level must be within volume data range.")
    if rc != 0:
        raise MemoryError(f"hvo_marching_cubes failed ({rc})")
    try:
        nv, nf = int(L.hvo_mesh_nv(handle)), int(L.hvo_mesh_nf(handle))
        if nv == 0:
            raise RuntimeError("No surface found at the given iso value.")
        verts = np.empty((nv, 3), np.float32)
        normals = np.empty((nv, 3), np.float32)
        keys = np.empty(nv, np.int64)
        faces = np.empty((nf, 3), np.int32)
        cubes = np.empty(nf, np.int64)
        L.hvo_mesh_copy(handle, _p(verts), _p(normals), _p(keys), _p(faces), _p(cubes))
    finally:
        L.hvo_mc_free(handle)
    values = np.full(nv, np.float32(level))
    if with_cubes:
        return verts, faces, normals, values, keys, cubes
    return (verts, faces, normals, values, keys) if with_keys else (verts, faces, normals, values)


def product_order(keys):
    """Permutation that brings the oracle's vertices (numbered in first-touch order like skimage's) into the order the
    CUDA product emits: vertices on lattice edges sorted by key = (owner voxel in scan order, axis), then the vertices
    inside cubes (key axis 3) sorted by cube.  skimage's own numbering is not reproduced by either (App. C)."""
    keys = np.asarray(keys, dtype=np.int64)
    return np.argsort(keys + ((keys & 3) == 3).astype(np.int64) * (1 << 62), kind="stable")


def get_mesh(tsdf_vol, color_vol, voxel_size, vol_origin):
    """TSDFVolume.get_mesh (fusion.py:332-350): marching cubes at level 0, then the in-tree post-processing."""
    verts, faces, norms, _ = marching_cubes(tsdf_vol, 0.0)
    verts_w, colors = fo.mesh_postprocess(verts, color_vol, voxel_size, vol_origin)
    return verts_w, faces, norms, colors


def mc_face_corners():
    """Corner ids (Lewiner numbering) of the six cube faces in mc_oracle.c's own face order."""
    out = np.zeros((6, 4), np.int32)
    for f in range(6):
        lib().hvo_mc_face_corners(ctypes.c_int(f), _p(out[f]))
    return out


def mc_tiling(config, pmask, tunnel):
    """Triangles (edge ids, 12 = centre vertex) and sub-case label mc_oracle.c derives for a (configuration, face
    resolution, tunnel) triple; ``pmask`` bit f = "the positive corners of oracle face f are joined".  None if the
    triple is not valid (a tunnel where Lewiner's decision tree has no interior test)."""
    tris = np.zeros(48, np.uint8)
    label = ctypes.create_string_buffer(32)
    L = lib()
    L.hvo_mc_tiling.restype = ctypes.c_int
    n = L.hvo_mc_tiling(ctypes.c_int(int(config)), ctypes.c_int(int(pmask)), ctypes.c_int(int(bool(tunnel))), _p(tris), label)
    if n < 0:
        return None
    return tris[:3 * n].reshape(n, 3).copy(), label.value.decode()


def mc_cube(values):
    """Decision of mc_oracle.c for one cube from its eight corner values (Lewiner corner order, level already
    subtracted): (configuration, pmask, tunnel, triangles, label)."""
    v = np.ascontiguousarray(values, dtype=np.float64)
    assert v.shape == (8,)
    tris = np.zeros(48, np.uint8)
    label = ctypes.create_string_buffer(32)
    cfg, pm, tn = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    L = lib()
    L.hvo_mc_cube.restype = ctypes.c_int
    n = L.hvo_mc_cube(_p(v), ctypes.byref(cfg), ctypes.byref(pm), ctypes.byref(tn), _p(tris), label)
    assert n >= 0
    return cfg.value, pm.value, tn.value, tris[:3 * n].reshape(n, 3).copy(), label.value.decode()
