"""ORACLE tooling (build container only): how fast is the reference's own CPU path, and how do the restatements that
travel to the GPU box compare?  Writes profiles/r02_reference_speed.json, which bench.py quotes
(``port_vs_reference_speed``) next to the CPU baselines it measures on the box.

    python -m oracle.measure_reference_speed            # needs /root/reference (numba JIT excluded from the timings)

Workload: BASELINE config C1 (256^3 voxels, 640x480 synthetic frames), `room` and `dense` scenes -- the largest
config the reference's CPU mode fits comfortably (its (N, 3) int64 voxel table and float64 temporaries take
~200 B/voxel).
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from hive_b200.synthetic import Scene  # noqa: E402
from oracle import c_oracle, fusion_numba, ref_shim  # noqa: E402


def _time(vol, scene, frames, n):
    fr = frames[0]
    vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose, 1.0)  # JIT / page faults
    t0 = time.perf_counter()
    for i in range(n):
        fr = frames[(i + 1) % len(frames)]
        vol.integrate(fr.rgb, fr.depth, scene.K, fr.pose, 1.0)
    return float(np.prod(vol._vol_dim)) * n / (time.perf_counter() - t0)


def main():
    assert ref_shim.available(), "/root/reference is needed"
    ref = ref_shim.reference_fusion()
    c_oracle.build()
    cores = c_oracle.use_all_host_threads()
    out = {"cores": cores, "cpu": open("/proc/cpuinfo").read().split("model name")[1].split("\n")[0].strip(": \t"),
           "workload": "C1: 256^3 voxels, 640x480 synthetic frames", "unit": "voxel-updates/s", "scenes": {}}
    for variant in ("dense", "room"):
        scene = Scene("C1", variant)
        frames = [scene.frame(i) for i in range(4)]
        r = _time(ref.TSDFVolume(scene.vol_bnds.copy(), scene.voxel_size, use_gpu=False), scene, frames, 3)
        nb = _time(fusion_numba.NumbaTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size), scene, frames, 3)
        c = _time(c_oracle.COracleTSDFVolume(scene.vol_bnds.copy(), scene.voxel_size), scene, frames, 12)
        out["scenes"][variant] = {"reference_numpy_numba": r, "restatement_numba": nb, "c_port_openmp": c,
                                  "numba_restatement_vs_reference": nb / r, "c_port_vs_reference": c / r}
        print(variant, out["scenes"][variant])
    with open(os.path.join(ROOT, "profiles", "r02_reference_speed.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
