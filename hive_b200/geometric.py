"""GPU drop-ins for the back-projection functions of ``hive/geometric.py`` (lines 66-78, 107-206).

Same names, arguments, return shapes/dtypes and assertion behaviour as the reference; host numpy in, host numpy out,
the arithmetic (float64, like the reference) runs in hand-written sm_100a kernels behind the C ABI.  No CPU fallback.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from . import _xfer

__all__ = ["get_pose_components", "point_cloud_from_depth", "point_cloud_from_rgbd", "world2image", "image2world",
           "depth_to_mesh_batch", "DepthMeshBatch"]

_EYE = np.eye(3)
_ZERO_T = np.zeros((3, 1))


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def validate_shape(x, x_name, expected_shape):
    """hive/utils.py:38-63 (None = any size)."""
    assert len(x.shape) == len(expected_shape), \
        f"{x_name} was expected to have {len(expected_shape)} dimensions but got {len(x.shape)}."
    for dim, expected in zip(x.shape, expected_shape):
        assert expected is None or dim == expected, \
            f"{x_name} was expected to have the shape {expected_shape}, but got {x.shape}."


def validate_camera_parameter_shapes(K, R, t):
    validate_shape(K, 'K', (3, 3))
    validate_shape(R, 'R', (3, 3))
    validate_shape(t, 't', (3, 1))


def get_pose_components(pose):
    """hive/geometric.py:66-78: R (3,3) and t (3,1) views of a 4x4 pose."""
    validate_shape(pose, 'pose', (4, 4))
    return pose[:3, :3], pose[:3, 3:]


def _kinv64(K):
    """inv(K) computed in K's own dtype (float32 from HIVE, hive/io.py:1100), then widened: geometric.py:202."""
    return np.ascontiguousarray(np.linalg.inv(np.asarray(K)), dtype=np.float64)


def _cam21(M, R, t):
    return np.concatenate([np.asarray(M, dtype=np.float64).reshape(9), np.asarray(R, dtype=np.float64).reshape(9),
                           np.asarray(t, dtype=np.float64).reshape(3)])


class DepthMeshBatch:
    """Concatenated meshes of a batch of frames (device tensors) with per-frame offsets."""

    def __init__(self, verts, faces, points2d, vdepth, frame_offsets):
        self.verts, self.faces, self.points2d, self.vdepth = verts, faces, points2d, vdepth
        self.frame_offsets = frame_offsets  # host (B+1, 2) int64

    def __len__(self):
        return self.frame_offsets.shape[0] - 1

    def frame(self, i):
        """(vertices (N,3) f64, faces (F,3) i32, points2d (N,2) i32, depth (N,) f32) of frame i as host numpy."""
        v0, f0 = self.frame_offsets[i]
        v1, f1 = self.frame_offsets[i + 1]
        return _xfer.to_host(self.verts[v0:v1], self.faces[f0:f1],
                             self.points2d[v0:v1] if self.points2d is not None else None,
                             self.vdepth[v0:v1] if self.vdepth is not None else None)

    def to_host(self):
        """(verts, faces, points2d, vdepth) of the whole batch as host numpy (faces index per-frame vertex ranges)."""
        return _xfer.to_host(self.verts, self.faces, self.points2d, self.vdepth)


_TRACE = None   # tools/time_d2m_phases.py sets a list: (label, perf_counter) after every host-side step of the batched call
_D2M_SCRATCH = {}  # (device index, B, H, W, want_points, stream) -> (scratch, aligned pointer, size, totals pair)
_D2M_CAPS = {}  # (device index, B, H, W, want_points) -> (vertices, faces) of the last batch of that shape


def depth_to_mesh_device(depth, mask, kinv, cams, max_depth_distance, flags, diag=None, want_points=True,
                         vertex_dtype=np.float64):
    """Core batched call on device tensors.  depth (B,H,W) f32, mask (B,H,W) u8 or None, cams (B,12) f64.

    The output sizes are data dependent.  The first batch of a shape reads the totals between the count and the emit
    launch; later batches of that shape emit straight away into buffers sized 1.25 x the previous batch (consecutive batches
    of a clip are alike), so count, scan and emit run back to back and the host reads the totals and the frame offsets in
    one go afterwards -- if the guess was too small, the emit is repeated exactly sized.

    ``vertex_dtype=np.float32`` (HV_D2M_VERTS_F32) stores the float64 vertices rounded to float32 -- what a glTF export keeps
    anyway -- for callers that pay for the bytes (host transfers); the reference's arrays are float64, the default.
    """
    torch = _lib.require_cuda()
    lib = _lib.load()
    B, H, W = (int(s) for s in depth.shape)
    dev = depth.device
    f32 = np.dtype(vertex_dtype) == np.float32
    assert f32 or np.dtype(vertex_dtype) == np.float64, "vertices are float64 (the reference) or float32"
    if f32:
        flags = int(flags) | 4  # HV_D2M_VERTS_F32
    key = (dev.index, B, H, W, bool(want_points), f32)
    kinv = np.ascontiguousarray(kinv, dtype=np.float64)

    def mark(label):
        if _TRACE is not None:
            import time
            _TRACE.append((label, time.perf_counter()))

    mark("start")
    with torch.cuda.device(dev):
        stream_id = torch.cuda.current_stream().cuda_stream
        stream = ctypes.c_void_p(stream_id)
        # scratch and the totals pair are kept per (shape, stream): calls on one stream are ordered, so the next call may
        # overwrite them (everything a call hands out lives in its own tensors)
        kept = _D2M_SCRATCH.get(key + (stream_id,))
        if kept is None:
            nbytes = ctypes.c_int64(0)
            _lib.check(lib.hv_d2m_scratch_bytes(B, H, W, ctypes.byref(nbytes)))
            scratch = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=dev)
            kept = (scratch, ctypes.c_void_p(scratch.data_ptr() + (-scratch.data_ptr()) % 256), nbytes,
                    torch.empty(2, dtype=torch.int64, device=dev))
            if len(_D2M_SCRATCH) >= 8:
                _D2M_SCRATCH.clear()
            _D2M_SCRATCH[key + (stream_id,)] = kept
        scratch, sptr, nbytes, totals = kept
        mdd = ctypes.c_float(np.float32(max_depth_distance))
        _lib.check(lib.hv_d2m_count(_ptr(depth), _ptr(mask), _ptr(diag), B, H, W, mdd, flags, sptr, int(nbytes.value),
                                    _ptr(totals), stream))
        mark("scratch + count launched")

        def outputs(n_verts, n_faces):
            return (torch.empty((n_verts, 3), dtype=torch.float32 if f32 else torch.float64, device=dev),
                    torch.empty((n_faces, 3), dtype=torch.int32, device=dev),
                    torch.empty((n_verts, 2), dtype=torch.int32, device=dev) if want_points else None,
                    torch.empty((n_verts,), dtype=torch.float32, device=dev) if want_points else None,
                    # the frame offsets are stored by the emit kernel straight into page-locked host memory (the address is
                    # valid on the device: unified addressing) -- 16 bytes per frame over the bus as posted writes, no copy
                    # to queue behind a bulk download on the D2H engine; visible to the host after the stream's sync
                    torch.empty((B + 1, 2), dtype=torch.int64, pin_memory=True))

        guess = _D2M_CAPS.get(key)
        frame_off_host = None
        if guess is not None:
            cap_v, cap_f = (min(int(1.25 * g) + 4096, lim) for g, lim in zip(guess, (B * H * W, 8 * B * H * W)))
            verts, faces, points2d, vdepth, frame_off = outputs(cap_v, cap_f)
            _lib.check(lib.hv_d2m_emit_capped(_ptr(depth), _ptr(mask), _ptr(diag), B, H, W, mdd, flags,
                                              kinv.ctypes.data_as(ctypes.c_void_p), _ptr(cams), sptr, int(nbytes.value), cap_v,
                                              cap_f, _ptr(verts), _ptr(faces), _ptr(points2d), _ptr(vdepth), _ptr(frame_off),
                                              _ptr(totals), stream))
            mark("outputs + emit launched")
            torch.cuda.current_stream().synchronize()  # the one wait of the call
            frame_off_host = frame_off.numpy()             # frame offsets, the last row = the totals
            n_verts, n_faces = (int(v) for v in frame_off_host[B])
            mark("offsets read")
            if n_verts <= cap_v and n_faces <= cap_f:
                verts, faces = verts[:n_verts], faces[:n_faces]
                if want_points:
                    points2d, vdepth = points2d[:n_verts], vdepth[:n_verts]
            else:
                frame_off_host = None  # too small a guess: again below, exactly sized
        else:
            n_verts, n_faces = (int(v) for v in _xfer.to_host(totals))
        if frame_off_host is None:
            verts, faces, points2d, vdepth, frame_off = outputs(n_verts, n_faces)
            _lib.check(lib.hv_d2m_emit(_ptr(depth), _ptr(mask), _ptr(diag), B, H, W, mdd, flags,
                                       kinv.ctypes.data_as(ctypes.c_void_p), _ptr(cams), sptr, int(nbytes.value), n_verts,
                                       n_faces, _ptr(verts), _ptr(faces), _ptr(points2d), _ptr(vdepth), _ptr(frame_off), stream))
            torch.cuda.current_stream().synchronize()
            frame_off_host = frame_off.numpy()
        _D2M_CAPS[key] = (n_verts, n_faces)
        mark("done")
        return DepthMeshBatch(verts, faces, points2d, vdepth, frame_off_host)


def d2m_flags(max_pixel_distance) -> int:
    """The lattice rule is exact for max_pixel_distance in [sqrt2, sqrt5); HIVE's default is 2.  Below
    sqrt2 no triangle of lattice points survives the edge filter (the shortest one has a sqrt2 hypotenuse): vertices only."""
    mpd = float(max_pixel_distance)
    if mpd < np.sqrt(2.0):
        return 2  # HV_D2M_VERTS_ONLY
    if mpd < 2.0:
        return 0
    if mpd < np.sqrt(5.0):
        return 1  # HV_D2M_BRIDGES
    raise NotImplementedError(
        f"max_pixel_distance={mpd}: the closed-form lattice triangulation covers [0, sqrt(5)) only")


def depth_to_mesh_batch(depths, masks, K, poses, max_pixel_distance=2, max_depth_distance=0.1, diag=None):
    """Per-frame foreground meshes for a batch of frames (hive/pipeline.py:386-409 fused).

    depths (B,H,W) float32, masks (B,H,W) bool/uint8 or None, K (3,3), poses (B,4,4) world-to-camera
    (hive/pipeline.py:357).  Returns a ``DepthMeshBatch``.
    """
    torch = _lib.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device())
    depths = np.asarray(depths)
    poses = np.asarray(poses, dtype=np.float64)
    B = depths.shape[0]
    cams = np.concatenate([poses[:, :3, :3].reshape(B, 9), poses[:, :3, 3].reshape(B, 3)], axis=1)
    d = _xfer.to_device(depths, dev, np.float32)
    m = None if masks is None else _xfer.to_device(masks, dev, np.uint8)
    g = None if diag is None else _xfer.to_device(diag, dev, np.uint8)
    c = _xfer.to_device(cams, dev)
    return depth_to_mesh_device(d, m, _kinv64(K), c, max_depth_distance, d2m_flags(max_pixel_distance), g)


_HOST_STREAMS = {}  # device index -> (upload, kernel, download) streams of depth_to_mesh_host
_HOST_TRACE = None  # tools/tune_d2m_host.py sets a list: (chunk, step, perf_counter) of depth_to_mesh_host's main thread


def _host_mark(chunk, step):
    if _HOST_TRACE is not None:
        import time
        _HOST_TRACE.append((chunk, step, time.perf_counter()))


class HostDepthMeshes:
    """Per-frame meshes as HOST arrays, chunk by chunk (what ``depth_to_mesh_host`` returns): ``frame(i)`` is
    (vertices (N,3) f64, faces (F,3) i32 frame-local, points2d (N,2) i32, depth (N,) f32), views of pinned blocks."""

    def __init__(self, chunks, n_frames):
        self._chunks, self._n = chunks, n_frames  # chunks: (first frame, verts, faces, points2d, vdepth, frame_offsets)

    def __len__(self):
        return self._n

    def frame(self, i):
        for first, verts, faces, points2d, vdepth, off in self._chunks:
            k = i - first
            if 0 <= k < off.shape[0] - 1:
                (v0, f0), (v1, f1) = off[k], off[k + 1]
                return (verts[v0:v1], faces[f0:f1], None if points2d is None else points2d[v0:v1],
                        None if vdepth is None else vdepth[v0:v1])
        raise IndexError(i)

    @property
    def n_vertices(self):
        return sum(int(c[5][-1, 0]) for c in self._chunks)

    @property
    def n_faces(self):
        return sum(int(c[5][-1, 1]) for c in self._chunks)


def depth_to_mesh_host(depths, masks, K, poses, max_pixel_distance=2, max_depth_distance=0.1, chunk_frames=20,
                       copy_threads=8, vertex_dtype=np.float64, want_points=True, max_pinned_bytes=1 << 30):
    """``depth_to_mesh_batch`` for host arrays in and host arrays out, as a pipeline over chunks of frames: while chunk c
    is triangulated, chunk c + 1 is on its way up (pinned staging filled by ``copy_threads`` host threads, then H2D on its
    own stream) and chunk c - 1 on its way down (D2H on a third stream) -- PCIe carries 1.5 MB per frame up and ~2.8 MB
    down (float64 vertices), the kernels are 1 % of that.  Same per-frame results as ``depth_to_mesh_batch(...).frame(i)``.
    The bus is the limit, so the two options are about bytes: ``vertex_dtype=np.float32`` (12 instead of 24 bytes per vertex)
    and ``want_points=False`` (no (u, v, depth) per vertex: 12 bytes less; ``frame(i)`` then returns None for them).
    At most ``max_pinned_bytes`` of input are staged in page-locked memory at a time (longer clips run part after part).
    """
    from concurrent.futures import ThreadPoolExecutor
    torch = _lib.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device())
    depths = np.asarray(depths)
    masks = None if masks is None else np.asarray(masks)
    poses = np.asarray(poses, dtype=np.float64)
    B, H, W = depths.shape
    part = max(1, int(chunk_frames), int(max_pinned_bytes) // (5 * H * W))
    if B > part:  # a long clip: part after part, each staged on its own
        chunks = []
        for a in range(0, B, part):
            sub = depth_to_mesh_host(depths[a:a + part], None if masks is None else masks[a:a + part], K, poses[a:a + part],
                                     max_pixel_distance, max_depth_distance, chunk_frames, copy_threads, vertex_dtype,
                                     want_points, max_pinned_bytes)
            chunks.extend((c[0] + a,) + tuple(c[1:]) for c in sub._chunks)
        return HostDepthMeshes(chunks, B)
    flags, kinv = d2m_flags(max_pixel_distance), _kinv64(K)
    cams_host = np.concatenate([poses[:, :3, :3].reshape(B, 9), poses[:, :3, 3].reshape(B, 3)], axis=1)
    n = max(1, min(int(chunk_frames), B))
    bounds, a, step = [], 0, max(1, n // 4)  # the first chunks are small: the pipeline starts after 1/4 of a chunk is staged
    while a < B:
        bounds.append((a, min(a + step, B)))
        a += step
        step = min(n, 2 * step)
    slots = 3
    # the whole clip is staged into pinned memory by the host threads, chunk by chunk in order, as fast as they can; the
    # loop below only waits for the chunk it is about to send (staging chunk c + 2 while chunk c runs made the staging
    # threads the bottleneck: next to the two DMA streams they copy at 4-6 GB/s each)
    pin_d = torch.empty((B, H, W), dtype=torch.float32, pin_memory=True)
    pin_m = torch.empty((B, H, W), dtype=torch.uint8, pin_memory=True) if masks is not None else None
    dev_d = [torch.empty((n, H, W), dtype=torch.float32, device=dev) for _ in range(slots)]
    dev_m = [torch.empty((n, H, W), dtype=torch.uint8, device=dev) for _ in range(slots)] if masks is not None else None
    with torch.cuda.device(dev):
        cams = _xfer.to_device(cams_host, dev)
        main = torch.cuda.current_stream()
        # the same three streams for every call: torch's caching allocator keeps its blocks per stream, so fresh streams
        # would cudaMalloc the scratch and the outputs of the first chunks again on every call (measured: 5 ms)
        streams = _HOST_STREAMS.get(dev.index)
        if streams is None:
            streams = _HOST_STREAMS[dev.index] = (torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream())
        s_in, s_run, s_out = streams
        for s_ in streams:
            s_.wait_stream(main)
        slot_free = [None] * slots  # event: the kernels that read device slot k have run

        def stage(c):
            a, b = bounds[c]
            rows = np.array_split(np.arange(a, b), copy_threads)

            def copy(idx):
                _host_mark(c, "  worker: copy begins")
                if len(idx):
                    lo, hi = int(idx[0]), int(idx[-1]) + 1
                    np.copyto(pin_d.numpy()[lo:hi], depths[lo:hi], casting="unsafe")
                    if masks is not None:
                        np.copyto(pin_m.numpy()[lo:hi], masks[lo:hi], casting="unsafe")
                _host_mark(c, "  worker: copy ends")
            return [pool.submit(copy, idx) for idx in rows]

        def upload(c, futures):
            for f in futures:
                f.result()
            a, b = bounds[c]
            k = c % slots
            with torch.cuda.stream(s_in):
                if slot_free[k] is not None:
                    s_in.wait_event(slot_free[k])
                dev_d[k][:b - a].copy_(pin_d[a:b], non_blocking=True)
                if masks is not None:
                    dev_m[k][:b - a].copy_(pin_m[a:b], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(s_in)
            return ev

        chunks, downloads = [], []
        with ThreadPoolExecutor(max_workers=max(1, int(copy_threads))) as pool:
            staged = {c: stage(c) for c in range(len(bounds))}
            uploaded = {0: upload(0, staged.pop(0))}
            for c, (a, b) in enumerate(bounds):
                _host_mark(c, "begin")
                if c + 1 < len(bounds):
                    uploaded[c + 1] = upload(c + 1, staged.pop(c + 1))  # on its way up while chunk c is worked on
                _host_mark(c, "next upload queued")
                k = c % slots
                with torch.cuda.stream(s_run):
                    s_run.wait_event(uploaded.pop(c))
                    out = depth_to_mesh_device(dev_d[k][:b - a], None if masks is None else dev_m[k][:b - a], kinv,
                                               cams[a:b], max_depth_distance, flags, want_points=want_points,
                                               vertex_dtype=vertex_dtype)
                    done = torch.cuda.Event()
                    done.record(s_run)
                slot_free[k] = done
                _host_mark(c, "kernels done")
                with torch.cuda.stream(s_out):  # (depth_to_mesh_device has synchronised s_run: the results exist)
                    host = []
                    for t in (out.verts, out.faces, out.points2d, out.vdepth):
                        if t is None:
                            host.append(None)
                            continue
                        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
                        if t.numel():
                            h.copy_(t, non_blocking=True)
                            t.record_stream(s_out)
                        host.append(h)
                downloads.append(host)
                chunks.append((a, out.frame_offsets))
                _host_mark(c, "download queued")
            s_out.synchronize()
            _host_mark(len(bounds), "downloads done")
        main.wait_stream(s_run)
    return HostDepthMeshes([(a,) + tuple(None if h is None else h.numpy() for h in host) + (off,)
                            for (a, off), host in zip(chunks, downloads)], B)


def point_cloud_from_depth(depth, mask, K, R=_EYE, t=_ZERO_T):
    """hive/geometric.py:107-126: (N, 3) float64 world points of ``mask & (depth > 0)``, row-major order."""
    torch = _lib.require_cuda()
    validate_camera_parameter_shapes(K, R, t)
    dev = torch.device("cuda", torch.cuda.current_device())
    depth = np.asarray(depth)
    assert np.asarray(mask).shape == depth.shape
    d = _xfer.to_device(depth[None], dev, np.float32)
    m = _xfer.to_device(np.asarray(mask)[None], dev, np.uint8)
    cams = torch.from_numpy(_cam21(np.eye(3), R, t)[9:][None].copy()).to(dev)
    out = depth_to_mesh_device(d, m, _kinv64(K), cams, 0.0, 2, want_points=False)  # HV_D2M_VERTS_ONLY
    return _xfer.to_host(out.verts)


def point_cloud_from_rgbd(rgb, depth, mask, K, R=_EYE, t=_ZERO_T):
    """hive/geometric.py:129-152: points plus (N, 4) RGBA (alpha 255) in the dtype of ``rgb``."""
    points = point_cloud_from_depth(depth, mask, K, R, t)
    valid = np.asarray(mask) & (np.asarray(depth) > 0.0)
    colour = np.zeros(shape=(len(points), 4), dtype=rgb.dtype)
    colour[:, :3] = rgb[valid]
    colour[:, 3] = 255
    return points, colour


def image2world(points, depth, K, R=_EYE, t=_ZERO_T, scale_factor=1.0):
    """hive/geometric.py:183-206."""
    torch = _lib.require_cuda()
    points, depth = np.asarray(points), np.asarray(depth)
    validate_shape(points, 'points', (None, 2))
    validate_shape(depth, 'depth', (points.shape[0],))
    validate_camera_parameter_shapes(K, R, t)
    dev = torch.device("cuda", torch.cuda.current_device())
    n = points.shape[0]
    p = _xfer.to_device(points, dev, np.float64)
    d = _xfer.to_device(depth, dev, np.float64)
    cam = torch.from_numpy(_cam21(_kinv64(K), R, t)).to(dev)
    out = torch.empty((n, 3), dtype=torch.float64, device=dev)
    _lib.check(_lib.load().hv_image2world(_ptr(p), _ptr(d), n, _ptr(cam), float(scale_factor), _ptr(out),
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return _xfer.to_host(out)


def world2image(points, K, R=_EYE, t=_ZERO_T, scale_factor=1.0, dtype=np.int32):
    """hive/geometric.py:155-180: ((N, 2) pixel coordinates in ``dtype``, (N,) float64 depth)."""
    torch = _lib.require_cuda()
    points = np.asarray(points)
    validate_shape(points, 'points', (None, 3))
    validate_camera_parameter_shapes(K, R, t)
    dev = torch.device("cuda", torch.cuda.current_device())
    n = points.shape[0]
    p = _xfer.to_device(points, dev, np.float64)
    cam = torch.from_numpy(_cam21(np.asarray(K, dtype=np.float64), R, t)).to(dev)
    uv = torch.empty((n, 2), dtype=torch.float64, device=dev)
    depth = torch.empty((n,), dtype=torch.float64, device=dev)
    _lib.check(_lib.load().hv_world2image(_ptr(p), n, _ptr(cam), float(scale_factor),
                                          int(issubclass(dtype, np.integer)), _ptr(uv), _ptr(depth),
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    uv_h, depth_h = _xfer.to_host(uv, depth)
    return np.array(uv_h, dtype=dtype), depth_h
