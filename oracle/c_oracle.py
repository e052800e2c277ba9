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
    srcs = [os.path.join(_HERE, f) for f in ("tsdf_oracle.c", "mc_oracle.c", "mc_tables.h", "d2m_oracle.c")]
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
        H, W = depth_im.shape
        folded = np.ascontiguousarray(fo.fold_color(color_im), dtype=np.float32)
        depth = np.ascontiguousarray(depth_im, dtype=np.float32)
        w2c = np.ascontiguousarray(np.linalg.inv(np.asarray(cam_pose, dtype=np.float64))[:3, :])
        k = np.asarray(cam_intr).astype(np.float32)
        intr = np.array([k[0, 0], k[1, 1], k[0, 2], k[1, 2]], dtype=np.float32)
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
