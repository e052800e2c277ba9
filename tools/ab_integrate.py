#!/usr/bin/env python
"""A/B timing of the integrate kernels (include/hive_b200.h, bits 8..15 of `flags`) on the C2 workload.

    python tools/ab_integrate.py [--config C2] [--frames 16] [--timed 90] [--kernels 1,2,0]

For every scene variant and kernel: fresh volume, untimed warm-up, CUDA-event time over `--timed` launches with the
frames resident in HBM, then a checksum of the three planes (all kernels must agree bit for bit).
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--timed", type=int, default=90)
    ap.add_argument("--warm", type=int, default=10)
    ap.add_argument("--kernels", default="1,2,3")
    ap.add_argument("--scenes", default="dense,room")
    args = ap.parse_args()
    kernels = [int(k) for k in args.kernels.split(",")]
    frames = {v: bench.make_frames(args.config, v, args.frames) for v in args.scenes.split(",")}

    import torch
    from hive_b200.fusion import TSDFVolume
    peak, _ = bench.measured_peak()
    for variant, fr in frames.items():
        sc = bench._scene(args.config, variant)
        digests, tsdf_ref = {}, None
        for k in kernels:
            vol = TSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
            vol.integrate_kernel = k
            dev = vol.upload_frames(fr, sc.K)
            n = len(dev)
            for i in range(args.warm):
                vol.integrate_device(dev, i % n)
            vol.count_updates = True
            vol.reset_update_counter()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.timed):
                vol.integrate_device(dev, (args.warm + i) % n)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.timed
            upd = vol.num_updated() / args.timed
            nvox = int(np.prod(vol._local_dim))
            gbs = (24.0 * upd + 8.0 * sc.H * sc.W) / (ms * 1e-3) / 1e9
            h = hashlib.sha256()  # weight and colour planes must agree bit for bit, tsdf within the 1e-5 gate
            for plane in (vol._weight_vol_gpu, vol._color_vol_gpu):
                h.update(plane.cpu().numpy().tobytes())
            digests[k] = h.hexdigest()[:16]
            if tsdf_ref is None:
                tsdf_ref = vol._tsdf_vol_gpu.clone()
            tsdf_dev = float((vol._tsdf_vol_gpu - tsdf_ref).abs().max().item())
            assert tsdf_dev <= 1e-5, f"kernel {k} deviates from kernel {kernels[0]} by {tsdf_dev} on tsdf"
            print(json.dumps({"scene": variant, "kernel": k, "ms_per_frame": round(ms, 4), "updated_fraction": upd / nvox,
                              "achieved_gbs": round(gbs, 1), "frac_of_peak": round(gbs / peak, 4),
                              "gvox_per_s": round(nvox / ms / 1e6, 1), "sha": digests[k], "tsdf_dev": tsdf_dev}), flush=True)
            del vol, dev
            torch.cuda.empty_cache()
        assert len(set(digests.values())) == 1, f"kernels disagree on {variant}: {digests}"
    print("all kernels agree (weight, colour bit-identical; tsdf within 1e-5)")


if __name__ == "__main__":
    main()
