"""ORACLE -- test infrastructure, not product code.

The reference's CPU path with the reference's own *shape*: ``third_party/tsdf_fusion_python/fusion.py:166-215,
260-303`` keeps an (N, 3) int64 table of every voxel coordinate, runs three numba ``prange`` kernels over all N voxels
and moves the rest through whole-volume numpy temporaries and fancy indexing.  ``oracle/fusion_oracle.py`` restates the
arithmetic slab by slab and ``oracle/tsdf_oracle.c`` as one fused C loop -- both much faster than what a HIVE user
without a GPU actually runs.  This module keeps the temporaries, the kernels and the gather/scatter pattern, so that
its *speed* stands for the reference's (bench.py ``--impl reference`` uses it when numba imports on the box;
``oracle/measure_reference_speed.py`` records in the build container how close it is to the real thing).

Pinned like the other restatements: planes bit-equal to the golden vectors dumped from the reference itself
(tests/test_oracle_tsdf.py) and to the live reference where /root/reference exists.
Only ``tests/`` and ``bench.py``'s CPU legs may import it.
"""
from __future__ import annotations

import numpy as np
from numba import njit, prange

from . import fusion_oracle as fo


@njit(parallel=True)
def _world_points(origin, coords, voxel_size):  # fusion.py:178-189
    origin = origin.astype(np.float32)
    coords32 = coords.astype(np.float32)
    out = np.empty_like(coords32, dtype=np.float32)
    for i in prange(coords32.shape[0]):
        for j in range(3):
            out[i, j] = origin[j] + (voxel_size * coords32[i, j])
    return out


@njit(parallel=True)
def _pixels(cam_pts, intr):  # fusion.py:191-203
    intr = intr.astype(np.float32)
    fx, fy, cx, cy = intr[0, 0], intr[1, 1], intr[0, 2], intr[1, 2]
    pix = np.empty((cam_pts.shape[0], 2), dtype=np.int64)
    for i in prange(cam_pts.shape[0]):
        pix[i, 0] = int(np.round((cam_pts[i, 0] * fx / cam_pts[i, 2]) + cx))
        pix[i, 1] = int(np.round((cam_pts[i, 1] * fy / cam_pts[i, 2]) + cy))
    return pix


@njit(parallel=True)
def _running_mean(tsdf, dist, w_old, obs_weight):  # fusion.py:205-215
    t_new = np.empty_like(tsdf, dtype=np.float32)
    w_new = np.empty_like(w_old, dtype=np.float32)
    for i in prange(len(tsdf)):
        w_new[i] = w_old[i] + obs_weight
        t_new[i] = (w_old[i] * tsdf[i] + obs_weight * dist[i]) / w_new[i]
    return t_new, w_new


class NumbaTSDFVolume:
    """``TSDFVolume(vol_bnds, voxel_size, use_gpu=False)`` of the reference, same memory and loop structure."""

    def __init__(self, vol_bnds, voxel_size):
        self._vol_bnds, self._vol_dim, self._vol_origin = fo.volume_geometry(vol_bnds, voxel_size)
        self._voxel_size = float(voxel_size)
        self._trunc_margin = 5 * self._voxel_size
        self._tsdf_vol_cpu = np.ones(self._vol_dim, dtype=np.float32)
        self._weight_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self._color_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        grids = np.meshgrid(*(range(int(d)) for d in self._vol_dim), indexing="ij")  # fusion.py:166-176
        self.vox_coords = np.concatenate([g.reshape(1, -1) for g in grids], axis=0).astype(int).T
        self.last_updated = 0

    def integrate(self, color_im, depth_im, cam_intr, cam_pose, obs_weight=1.0):
        im_h, im_w = depth_im.shape
        color = fo.fold_color(color_im)
        pts = _world_points(self._vol_origin, self.vox_coords, self._voxel_size)
        pts = fo.rigid_transform(pts, np.linalg.inv(cam_pose))  # (N, 4) float64 hstack + GEMM, fusion.py:353-358
        z = pts[:, 2]
        pix = _pixels(pts, cam_intr)
        u, v = pix[:, 0], pix[:, 1]
        in_view = (u >= 0) & (u < im_w) & (v >= 0) & (v < im_h) & (z > 0)
        d = np.zeros(u.shape)
        d[in_view] = depth_im[v[in_view], u[in_view]]
        diff = d - z
        hit = (d > 0) & (diff >= -self._trunc_margin)
        dist = np.minimum(1, diff / self._trunc_margin)
        ix, iy, iz = self.vox_coords[hit, 0], self.vox_coords[hit, 1], self.vox_coords[hit, 2]
        w_old = self._weight_vol_cpu[ix, iy, iz]
        t_new, w_new = _running_mean(self._tsdf_vol_cpu[ix, iy, iz], dist[hit], w_old, obs_weight)
        self._weight_vol_cpu[ix, iy, iz] = w_new
        self._tsdf_vol_cpu[ix, iy, iz] = t_new
        o_r, o_g, o_b = fo.unpack_color(self._color_vol_cpu[ix, iy, iz])
        n_r, n_g, n_b = fo.unpack_color(color[v[hit], u[hit]])
        mix = [np.minimum(255., np.round((w_old * o + obs_weight * n) / w_new)) for o, n in ((o_b, n_b), (o_g, n_g), (o_r, n_r))]
        self._color_vol_cpu[ix, iy, iz] = mix[0] * fo.COLOR_CONST + mix[1] * 256 + mix[2]
        self.last_updated = int(hit.sum())

    def get_volume(self):
        return self._tsdf_vol_cpu, self._color_vol_cpu
