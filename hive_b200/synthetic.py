"""Seeded synthetic RGB-D scenes for the parity tests, ``smoke()`` and ``bench.py``.

There is no network and no dataset on the GPU box, so every input of the hot path is generated here, on the
host, with numpy only.  The generator mimics what HIVE feeds the path:

* depth is ray-cast analytically, quantised to uint16 millimetres and converted exactly as
  ``hive/io.py:1032-1039`` does (``f32(0.001) * f32(mm)``, values above ``max_depth`` zeroed);
* colour is uint8 (H, W, 3) as PIL would deliver it (``hive/io.py:556-564``);
* intrinsics are float32 3x3 (``hive/io.py:1100``), Kinect/TUM values from ``hive/sensor.py:26``;
* poses are float64 4x4 camera-to-world built through float32 quaternions like ``hive/geometric.py:580-590``;
* the instance mask is uint8 (0 = background, 1..k = dynamic objects) as in ``hive/pipeline.py:340-371``.

Two scene variants:

``room``   camera inside a 5 x 3 x 5 m room with three spheres and one moving ellipsoid ("person");
           a realistic fraction of the voxels is updated per frame.
``dense``  camera outside the volume looking through it at a far wall, so (nearly) every voxel is in
           front of the surface and is updated every frame -- the HBM-roofline configuration.
"""
from __future__ import annotations

import dataclasses

import numpy as np

SEED = 20261104

#: BASELINE.json configs (C1..C5); ``dim`` voxels per axis, ``vs`` voxel size in metres.
CONFIGS = {
    "tiny": dict(dim=(40, 36, 46), vs=0.12, frames=6, H=60, W=80),
    "small": dict(dim=(96, 80, 104), vs=0.05, frames=8, H=120, W=160),
    "C1": dict(dim=(256, 256, 256), vs=0.02, frames=100, H=480, W=640),
    "C2": dict(dim=(512, 512, 512), vs=0.01, frames=300, H=480, W=640),
    "C3": dict(dim=(1024, 1024, 1024), vs=0.005, frames=64, H=1440, W=1920),
    "C5": dict(dim=(2048, 2048, 2048), vs=0.0025, frames=32, H=1080, W=1920),
}

ROOM_HALF = np.array([2.5, 1.5, 2.5])
ROOM_CENTRE = np.array([0.0, 0.0, 2.5])


def _rotation(rx: float, ry: float, rz: float) -> np.ndarray:
    """Rotation matrix from small Euler angles, passed through a float32 quaternion like the reference's
    Trajectory (7-vectors are float32 on disk; ``hive/geometric.py:363,580-590``)."""
    cx, sx = np.cos(rx / 2), np.sin(rx / 2)
    cy, sy = np.cos(ry / 2), np.sin(ry / 2)
    cz, sz = np.cos(rz / 2), np.sin(rz / 2)
    # q = qz * qy * qx  (x, y, z, w)
    q = np.array([
        sx * cy * cz - cx * sy * sz,
        cx * sy * cz + sx * cy * sz,
        cx * cy * sz - sx * sy * cz,
        cx * cy * cz + sx * sy * sz,
    ]).astype(np.float32).astype(np.float64)
    q /= np.linalg.norm(q)
    x, y, z, w = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
    ])


def _hash_u32(a: np.ndarray) -> np.ndarray:
    a = a.astype(np.uint32)
    a ^= a >> np.uint32(16)
    a *= np.uint32(0x7FEB352D)
    a ^= a >> np.uint32(15)
    a *= np.uint32(0x846CA68B)
    a ^= a >> np.uint32(16)
    return a


@dataclasses.dataclass
class Frame:
    rgb: np.ndarray  # (H, W, 3) uint8
    depth: np.ndarray  # (H, W) float32 metres, 0 = invalid
    mask: np.ndarray  # (H, W) uint8 instance ids
    pose: np.ndarray  # (4, 4) float64 camera-to-world


class Scene:
    """A seeded analytic scene; ``frame(i)`` is a pure function of (config, variant, seed, i)."""

    def __init__(self, config: str = "C1", variant: str = "room", seed: int = SEED, frames: int | None = None,
                 invalid_fraction: float = 0.03):
        cfg = dict(CONFIGS[config])
        self.config, self.variant, self.seed = config, variant, seed
        self.H, self.W = cfg["H"], cfg["W"]
        self.num_frames = frames if frames is not None else cfg["frames"]
        self.voxel_size = float(cfg["vs"])
        self.dim = tuple(int(d) for d in cfg["dim"])
        self.invalid_fraction = invalid_fraction
        # dense: the far wall sits behind the whole volume; 8 weak-scaling slabs of 512 x 1 cm put it ~47 m away, so
        # only the uint16-millimetre range itself (65.535 m) limits the depth of this variant
        self.max_depth = 10.0 if variant == "room" else 65.0

        # Kinect/TUM intrinsics scaled to the frame size (hive/sensor.py:26); float32 like hive/io.py:1100.
        s = self.W / 640.0
        if (self.H, self.W) == (1080, 1920):  # Unreal-style: 90 degree horizontal field of view
            f = self.W / 2.0
            K = [[f, 0, (self.W - 1) / 2.0], [0, f, (self.H - 1) / 2.0], [0, 0, 1]]
        else:
            K = [[580.0 * s, 0, 320.0 * s - 0.5], [0, 580.0 * s, 240.0 * s - 0.5], [0, 0, 1]]
        self.K = np.array(K, dtype=np.float32)

        # Volume: centred on the room; the upper bound sits half a voxel short so ceil() yields `dim` exactly.
        ext = np.array(self.dim) * self.voxel_size
        lo = ROOM_CENTRE - ext / 2.0
        hi = lo + (np.array(self.dim) - 0.5) * self.voxel_size
        self.vol_bnds = np.stack([lo, hi], axis=1)
        self.extent = ext

        rng = np.random.default_rng(seed)
        self._phase = rng.uniform(0, 2 * np.pi, size=6)
        self.spheres = [
            (np.array([-1.2, 0.9, 3.4]), 0.6),
            (np.array([1.3, 1.0, 2.8]), 0.5),
            (np.array([0.2, -0.6, 4.1]), 0.45),
        ]

    # ------------------------------------------------------------------ poses
    def pose(self, i: int) -> np.ndarray:
        """Camera-to-world 4x4 float64, walking_xyz-like: +-0.4 m sinusoids, rotation <= 5 degrees."""
        t = i / 30.0
        ph = self._phase
        trans = 0.4 * np.array([np.sin(0.9 * t + ph[0]), 0.5 * np.sin(1.3 * t + ph[1]), np.sin(0.7 * t + ph[2])])
        ang = np.deg2rad(5.0) * np.array([np.sin(1.1 * t + ph[3]), np.sin(0.8 * t + ph[4]), np.sin(0.5 * t + ph[5])])
        if self.variant == "room":
            centre = np.array([0.0, 0.0, 0.9])
        else:  # dense: far enough back that the frustum covers the whole near face of the volume
            half = 0.5 * max(self.extent[0] / (self.W / 2.0 / self.K[0, 0]), self.extent[1] / (self.H / 2.0 / self.K[1, 1]))
            centre = np.array([0.0, 0.0, ROOM_CENTRE[2] - self.extent[2] / 2.0 - 1.08 * half])
            trans = trans * 0.05
            ang = ang * 0.05
        M = np.eye(4)
        M[:3, :3] = _rotation(*ang)
        M[:3, 3] = (centre + trans).astype(np.float32)
        return M

    def person(self, i: int):
        """Moving ellipsoid: centre and semi-axes."""
        t = i / 30.0
        c = np.array([0.8 * np.sin(0.6 * t + 1.0), 0.55, 2.2 + 0.5 * np.cos(0.4 * t)])
        return c, np.array([0.28, 0.85, 0.25])

    # ------------------------------------------------------------------ frames
    def frame(self, i: int) -> Frame:
        H, W = self.H, self.W
        pose = self.pose(i)
        R, o = pose[:3, :3], pose[:3, 3]
        fx, fy, cx, cy = (float(self.K[0, 0]), float(self.K[1, 1]), float(self.K[0, 2]), float(self.K[1, 2]))
        u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
        dc = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], axis=-1).reshape(-1, 3)
        d = dc @ R.T  # world direction with camera-z component 1 => ray parameter == depth
        n = d.shape[0]
        best = np.full(n, np.inf)
        obj = np.zeros(n, dtype=np.int32)  # 0 walls, 1..3 spheres, 4 person, 5 far wall

        if self.variant == "room":
            lo, hi = ROOM_CENTRE - ROOM_HALF, ROOM_CENTRE + ROOM_HALF
            with np.errstate(divide="ignore", invalid="ignore"):
                for ax in range(3):
                    for plane in (lo[ax], hi[ax]):
                        tt = (plane - o[ax]) / d[:, ax]
                        ok = (tt > 1e-6) & (tt < best)
                        best = np.where(ok, tt, best)
        else:
            far_z = ROOM_CENTRE[2] + self.extent[2] / 2.0 + 0.35
            tt = (far_z - o[2]) / d[:, 2]
            ok = tt > 1e-6
            best = np.where(ok, tt, best)
            obj[:] = 5

        def quadric(centre, radii, ident):
            nonlocal best, obj
            oc = (o - centre) / radii
            dd = d / radii
            a = np.einsum("ij,ij->i", dd, dd)
            b = 2.0 * dd @ oc
            c = oc @ oc - 1.0
            disc = b * b - 4 * a * c
            hit = disc > 0
            sq = np.sqrt(np.where(hit, disc, 0.0))
            tt = (-b - sq) / (2 * a)
            ok = hit & (tt > 1e-6) & (tt < best)
            best = np.where(ok, tt, best)
            obj = np.where(ok, ident, obj)

        for k, (c, r) in enumerate(self.spheres):
            quadric(c, np.array([r, r, r]), k + 1)
        pc, pr = self.person(i)
        quadric(pc, pr, 4)

        depth_m = np.where(np.isfinite(best), best, 0.0)
        # uint16 millimetres on disk -> float32 metres exactly as hive/io.py:1032-1039
        mm = np.clip(np.rint(depth_m * 1000.0), 0, 65535).astype(np.uint16)
        depth = np.float32(1.0 / 1000.0) * mm.astype(np.float32)
        depth[depth > self.max_depth] = 0.0
        pix = np.arange(n, dtype=np.uint32)
        noise = _hash_u32(pix * np.uint32(2654435761) + np.uint32((self.seed + 7919 * i) & 0xFFFFFFFF))
        drop = (noise & np.uint32(0xFFFF)).astype(np.float64) / 65536.0 < self.invalid_fraction
        depth[drop] = 0.0

        # procedural colour: 25 cm checker keyed on the hit position, tinted per object, plus hash noise
        p = o[None, :] + d * depth_m[:, None]
        cell = np.floor(p / 0.25).astype(np.int64)
        chk = ((cell[:, 0] + cell[:, 1] + cell[:, 2]) & 1).astype(np.int32)
        base = np.array([[200, 190, 170], [220, 60, 50], [60, 200, 80], [70, 90, 220], [230, 200, 40], [150, 150, 160]])
        col = base[obj] - 60 * chk[:, None]
        col = col + ((noise >> np.uint32(16)) & np.uint32(31)).astype(np.int32)[:, None] - 16
        rgb = np.clip(col, 0, 255).astype(np.uint8).reshape(H, W, 3)

        mask = np.where(obj == 4, 1, 0).astype(np.uint8)
        mask = np.where(obj == 1, 2, mask).astype(np.uint8).reshape(H, W)  # a second "instance"
        return Frame(rgb=rgb, depth=depth.reshape(H, W), mask=mask, pose=pose)

    def frames(self, n: int | None = None):
        for i in range(self.num_frames if n is None else n):
            yield self.frame(i)


class _Trajectory:
    """The two calls the TSDF driver makes on ``HiveDataset.camera_trajectory`` (hive/fusion.py:51, 111)."""

    def __init__(self, matrices):
        self._m = np.asarray(matrices, dtype=np.float64)

    def inverse(self):
        return _Trajectory(np.linalg.inv(self._m))

    def to_homogenous_transforms(self):
        return self._m.copy()


class _Indexed:
    def __init__(self, fn):
        self._fn = fn

    def __getitem__(self, i):
        return self._fn(i)  # a fresh array per access, like decoding a PNG (hive/io.py:550)


def panning_poses(scene: "Scene", n: int, step_deg: float = 9.0):
    """World-to-camera matrices of the scene's trajectory with an extra pan of ``step_deg`` per frame about the camera's
    y axis, so that the overlap with earlier frames really falls (the stock trajectory barely moves): the trajectory the
    key-frame selection tests and goldens run on."""
    def rot_y(a):
        c, s = np.cos(a), np.sin(a)
        m = np.eye(4)
        m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s, -s, c
        return m
    return [np.linalg.inv(scene.pose(i) @ rot_y(np.deg2rad(step_deg * i))) for i in range(n)]


class SyntheticDataset:
    """Duck-typed stand-in for ``hive.io.HiveDataset`` as read by ``hive.fusion.tsdf_fusion`` (hive/fusion.py:79-134):
    world-to-camera trajectory, RGB / depth / instance-mask frame accessors, float32 camera matrix."""

    def __init__(self, scene: Scene, num_frames: int | None = None, has_inpainted_frame_data: bool = False):
        self.scene = scene
        self.num_frames = scene.num_frames if num_frames is None else int(num_frames)
        self.camera_matrix = scene.K
        self.has_inpainted_frame_data = has_inpainted_frame_data
        self._cache = {}
        self.camera_trajectory = _Trajectory([np.linalg.inv(scene.pose(i)) for i in range(self.num_frames)])
        self.bg_rgb_dataset = _Indexed(lambda i: self._frame(i).rgb.copy())
        self.bg_depth_dataset = _Indexed(lambda i: self._frame(i).depth.copy())
        self.mask_dataset = _Indexed(lambda i: self._frame(i).mask.copy())

    def _frame(self, i):
        if i not in self._cache:
            self._cache[i] = self.scene.frame(i)
        return self._cache[i]
