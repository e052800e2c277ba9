"""ORACLE -- test infrastructure, not product code.

CPU restatement (numpy) of the reference's TSDF integration, CPU mode:
``third_party/tsdf_fusion_python/fusion.py:25-57`` (constructor), ``:217-232`` + ``:260-303`` (integrate),
``:178-215`` (numba helpers ``vox2world`` / ``cam2pix`` / ``integrate_tsdf``), ``:353-358`` (``rigid_transform``),
``:361-375`` (``get_view_frustum``) and the post-marching-cubes lines ``:339-348`` of ``get_mesh``.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``
may import this module.  The product (``hive_b200``) never does.

Parity pinning: the reference publishes no test for this path.  This restatement is pinned against
the reference *itself*, run in the build container through ``oracle/ref_shim.py``:
``tests/test_oracle_vs_reference.py`` (live, skipped where /root/reference is absent) and the committed dumps
under ``tests/golden/`` produced by ``oracle/make_golden.py``.

Arithmetic contract restated here (each item verified against the reference's output):
  A.1 colour fold          floor(B*65536 + G*256 + R) in float32                         (fusion.py:231-232)
  A.2 world point          float32(float64(origin_j) + voxel_size * idx_j)                 (fusion.py:183-188)
  A.3 camera point         inv(cam_pose)[:3] @ [p, 1] in float64                           (fusion.py:263,356-358)
  A.4 pixel                int(round_half_even((q_x * fx) / q_z + cx)), fx.. float32->f64   (fusion.py:196-202)
  A.5 validity             0<=px<W, 0<=py<H, q_z>0, d>0, d-q_z >= -trunc                   (fusion.py:269-279)
  A.6 tsdf / weight        w' = f32(w+ow); tsdf' = f32((f64(f32(w*tsdf)) + ow*dist)/f64(w')) (fusion.py:207-214)
  A.7 colour               per channel min(255, round_half_even((w*old + ow*new)/w')) in f32 (fusion.py:292-303)

The volume is processed in x-slabs so that the temporaries stay small (the reference materialises ~200 B/voxel).
"""
from __future__ import annotations

import numpy as np

COLOR_CONST = np.float32(256 * 256)


def volume_geometry(vol_bnds, voxel_size):
    """Constructor arithmetic of fusion.py:33-46.  Returns (vol_bnds snapped, vol_dim int64[3], origin f32[3])."""
    vol_bnds = np.array(vol_bnds, dtype=np.float64)
    assert vol_bnds.shape == (3, 2), "[!] `vol_bnds` should be of shape (3, 2)."
    voxel_size = float(voxel_size)
    vol_dim = np.ceil((vol_bnds[:, 1] - vol_bnds[:, 0]) / voxel_size).copy(order="C").astype(int)
    vol_bnds[:, 1] = vol_bnds[:, 0] + vol_dim * voxel_size
    origin = vol_bnds[:, 0].copy(order="C").astype(np.float32)
    return vol_bnds, vol_dim, origin


def world_tables(origin_f32, voxel_size, vol_dim):
    """A.2: per-axis world coordinate of every voxel index, float64 arithmetic rounded to float32."""
    return [
        (origin_f32[j].astype(np.float64) + float(voxel_size) * np.arange(int(vol_dim[j]), dtype=np.float64)).astype(
            np.float32)
        for j in range(3)
    ]


def fold_color(color_im):
    """A.1 (fusion.py:231-232)."""
    c = np.asarray(color_im).astype(np.float32)
    return np.floor(c[..., 2] * COLOR_CONST + c[..., 1] * np.float32(256) + c[..., 0])


def unpack_color(packed_f32):
    """fusion.py:293-295 / :344-346: (r, g, b) float32 from the packed float32 colour."""
    b = np.floor(packed_f32 / COLOR_CONST)
    g = np.floor((packed_f32 - b * COLOR_CONST) / np.float32(256))
    r = packed_f32 - b * COLOR_CONST - g * np.float32(256)
    return r, g, b


class OracleTSDFVolume:
    """numpy restatement of the reference's ``TSDFVolume`` in CPU mode (fusion.py:21-303)."""

    def __init__(self, vol_bnds, voxel_size, slab=16):
        self._vol_bnds, self._vol_dim, self._vol_origin = volume_geometry(vol_bnds, voxel_size)
        self._voxel_size = float(voxel_size)
        self._trunc_margin = 5 * self._voxel_size
        self._tsdf_vol_cpu = np.ones(self._vol_dim, dtype=np.float32)
        self._weight_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self._color_vol_cpu = np.zeros(self._vol_dim, dtype=np.float32)
        self._tables = world_tables(self._vol_origin, self._voxel_size, self._vol_dim)
        self._slab = slab
        self.last_updated = 0  # U_f of the most recent integrate (valid_pts.sum(), fusion.py:279)

    def integrate(self, color_im, depth_im, cam_intr, cam_pose, obs_weight=1.0):
        im_h, im_w = depth_im.shape
        color = fold_color(color_im)
        depth_im = np.asarray(depth_im)
        T = np.linalg.inv(np.asarray(cam_pose, dtype=np.float64))  # fusion.py:263
        intr = np.asarray(cam_intr).astype(np.float32)
        fx, fy, cx, cy = (np.float64(intr[0, 0]), np.float64(intr[1, 1]), np.float64(intr[0, 2]), np.float64(intr[1, 2]))
        wx, wy, wz = (t.astype(np.float64) for t in self._tables)
        X, Y, Z = (int(d) for d in self._vol_dim)
        trunc = self._trunc_margin
        ow = float(obs_weight)
        ow32 = np.float32(obs_weight)
        updated = 0
        for x0 in range(0, X, self._slab):
            x1 = min(X, x0 + self._slab)
            px = wx[x0:x1, None, None]
            py = wy[None, :, None]
            pz = wz[None, None, :]
            # A.3 -- plain left-to-right float64 sums (the reference uses a BLAS dgemm; SURVEY.md §7 "Precision")
            qx = T[0, 0] * px + T[0, 1] * py + T[0, 2] * pz + T[0, 3]
            qy = T[1, 0] * px + T[1, 1] * py + T[1, 2] * pz + T[1, 3]
            qz = T[2, 0] * px + T[2, 1] * py + T[2, 2] * pz + T[2, 3]
            with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
                u = np.round((qx * fx) / qz + cx)  # A.4, half-to-even
                v = np.round((qy * fy) / qz + cy)
                inside = (u >= 0) & (u < im_w) & (v >= 0) & (v < im_h) & (qz > 0)  # A.5
            ui = np.where(inside, u, 0).astype(np.int64)
            vi = np.where(inside, v, 0).astype(np.int64)
            d = np.where(inside, depth_im[vi, ui].astype(np.float64), 0.0)
            diff = d - qz
            valid = (d > 0) & (diff >= -trunc)
            if not valid.any():
                continue
            updated += int(valid.sum())
            dist = np.minimum(1.0, diff[valid] / trunc)
            sl = (slice(x0, x1),)
            w_old = self._weight_vol_cpu[sl][valid]
            t_old = self._tsdf_vol_cpu[sl][valid]
            w_new = (w_old.astype(np.float64) + ow).astype(np.float32)  # A.6
            t_new = (((w_old * t_old).astype(np.float64) + ow * dist) / w_new.astype(np.float64)).astype(np.float32)
            tv = self._tsdf_vol_cpu[sl]
            wv = self._weight_vol_cpu[sl]
            cv = self._color_vol_cpu[sl]
            tv[valid] = t_new
            wv[valid] = w_new
            # A.7 -- everything float32
            o_r, o_g, o_b = unpack_color(cv[valid])
            n_r, n_g, n_b = unpack_color(color[vi[valid], ui[valid]])
            m255 = np.float32(255.0)
            b = np.minimum(m255, np.round((w_old * o_b + ow32 * n_b) / w_new))
            g = np.minimum(m255, np.round((w_old * o_g + ow32 * n_g) / w_new))
            r = np.minimum(m255, np.round((w_old * o_r + ow32 * n_r) / w_new))
            cv[valid] = b * COLOR_CONST + g * np.float32(256) + r
        self.last_updated = updated

    def get_volume(self):
        return self._tsdf_vol_cpu, self._color_vol_cpu


def rigid_transform(xyz, transform):
    """fusion.py:353-358."""
    xyz_h = np.hstack([xyz, np.ones((len(xyz), 1), dtype=np.float32)])
    return (transform @ xyz_h.T).T[:, :3]


def get_view_frustum(depth_im, cam_intr, cam_pose):
    """fusion.py:361-375: apex + four image corners pushed to the maximum depth, in world coordinates."""
    im_h, im_w = depth_im.shape[0], depth_im.shape[1]
    md = np.max(depth_im)
    zs = np.array([0, md, md, md, md])
    pts = np.array([
        (np.array([0, 0, 0, im_w, im_w]) - cam_intr[0, 2]) * zs / cam_intr[0, 0],
        (np.array([0, 0, im_h, 0, im_h]) - cam_intr[1, 2]) * zs / cam_intr[1, 1],
        zs,
    ])
    return rigid_transform(pts.T, cam_pose).T


def mesh_postprocess(verts_vox, color_vol, voxel_size, vol_origin):
    """In-tree part of get_mesh after marching cubes (fusion.py:339-348): world verts (f32) and uint8 colours."""
    verts_vox = np.asarray(verts_vox, dtype=np.float32)
    ind = np.round(verts_vox).astype(int)
    verts = verts_vox * voxel_size + vol_origin
    rgb = color_vol[ind[:, 0], ind[:, 1], ind[:, 2]]
    r, g, b = unpack_color(rgb)
    colors = np.floor(np.asarray([r, g, b])).T.astype(np.uint8)
    return verts, colors


def meshwrite(filename, verts, faces, norms, colors):
    """fusion.py:378-411, line for line in behaviour: header, one '%'-formatted row per vertex, one per face."""
    with open(filename, 'w') as f:
        f.write("ply\nformat ascii 1.0\n")
        f.write("element vertex %d\n" % (verts.shape[0]))
        for name in ("x", "y", "z", "nx", "ny", "nz"):
            f.write("property float %s\n" % name)
        for name in ("red", "green", "blue"):
            f.write("property uchar %s\n" % name)
        f.write("element face %d\n" % (faces.shape[0]))
        f.write("property list uchar int vertex_index\nend_header\n")
        for i in range(verts.shape[0]):
            f.write("%f %f %f %f %f %f %d %d %d\n" % (verts[i, 0], verts[i, 1], verts[i, 2], norms[i, 0], norms[i, 1],
                                                      norms[i, 2], colors[i, 0], colors[i, 1], colors[i, 2]))
        for i in range(faces.shape[0]):
            f.write("3 %d %d %d\n" % (faces[i, 0], faces[i, 1], faces[i, 2]))


def pcwrite(filename, xyzrgb):
    """fusion.py:414-438."""
    xyz = xyzrgb[:, :3]
    rgb = xyzrgb[:, 3:].astype(np.uint8)
    with open(filename, 'w') as f:
        f.write("ply\nformat ascii 1.0\n")
        f.write("element vertex %d\n" % (xyz.shape[0]))
        for name in ("x", "y", "z"):
            f.write("property float %s\n" % name)
        for name in ("red", "green", "blue"):
            f.write("property uchar %s\n" % name)
        f.write("end_header\n")
        for i in range(xyz.shape[0]):
            f.write("%f %f %f %d %d %d\n" % (xyz[i, 0], xyz[i, 1], xyz[i, 2], rgb[i, 0], rgb[i, 1], rgb[i, 2]))
