#!/usr/bin/env python
"""Benchmark of the TSDF hot path: BASELINE.json's metric on BASELINE.json's config.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scene dense|room]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...   (N > 1)

Workload (``config.workload``): BASELINE config C2 -- a 512^3 TSDF at 1 cm voxels integrating synthetic 640x480
RGB-D frames; one *step* integrates a batch of ``--batch`` (default 30) frames, so the default K=10 steps are the
300 frames of the config.  The headline scene is ``dense`` (camera outside the volume, U/N ~ 0.9: the HBM-roofline
configuration); the realistic ``room`` scene is measured too and reported under ``"room"``.

For N > 1 GPUs the volume is z-slab sharded (one 512^3 slab per GPU: weak scaling, the global volume is
512 x 512 x 512N), every frame batch is NCCL-broadcast from rank 0 and each rank integrates it into its slab.

``value``  : voxel-updates/s (voxels visited per frame x frames / time, as the reference README's FPS x volume),
             frames already resident in HBM as packed texels, CUDA-event timed, max over ranks.
``e2e``    : the same metric through the reference-facing call ``TSDFVolume.integrate(color_im, depth_im, K, pose)``
             with host numpy buffers: H2D of every frame and a D2H read of the update counter inside the timed region.
``roofline``: integrate kernel, algorithmic bytes (24 B per voxel the oracle updates + 8 B per pixel)
             over its CUDA-event launch duration, against MEASURED_PEAKS.json's HBM copy bandwidth.
``cpu_baseline``: the oracle's C/OpenMP port of the reference's CPU path (oracle/tsdf_oracle.c) on the host cores.

``parity_checked``: after the timed loops, random blocks of the volume the loops produced are compared with the oracle
             run over the same frame sequence (tsdf <= 1e-5, weight and colour bit-exact).
``build``   : hash compiled into the library that ran vs hash of the sources lying next to it.

``--impl reference`` times the reference's CPU path.  The reference is Python (numpy + numba): it cannot be compiled
into ``oracle/_ref`` and ``/root/reference`` does not travel, so the arm runs ``oracle/fusion_numba.py`` -- a restatement
with the reference's own memory and loop structure, measured at 0.99-1.06x the real reference's speed in the build
container (profiles/r02_reference_speed.json) -- when numba imports on the box, else the C/OpenMP port (42-69x faster
than the reference).  The line says which one ran.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

if "reference" in sys.argv:  # the reference arm uses every host thread (torchrun exports OMP_NUM_THREADS=1 to all ranks)
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMBA_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count() or 1)

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "tsdf_voxel_updates_per_s"
UNIT = "voxel-updates/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=30, help="frames per step")
    ap.add_argument("--config", default="C2")
    ap.add_argument("--scene", default="dense", choices=["dense", "room"])
    ap.add_argument("--distinct-frames", type=int, default=60, help="distinct synthetic frames generated (cycled)")
    ap.add_argument("--no-room", action="store_true", help="skip the secondary room-scene measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-mesh", action="store_true")
    ap.add_argument("--cpu-sample-frames", type=int, default=40)
    ap.add_argument("--d2m-frames", type=int, default=120, help="frames per depth-to-mesh batch")
    ap.add_argument("--no-d2m", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = one base volume per GPU along z (default); strong = the base volume cut in N slabs")
    ap.add_argument("--no-c5", action="store_true", help="skip the 2048^3 strong-scaling block (c5_strong)")
    ap.add_argument("--c5-frames", type=int, default=6, help="timed frames of the c5_strong block")
    ap.add_argument("--no-driver", action="store_true", help="skip the hive.fusion.tsdf_fusion driver measurement")
    ap.add_argument("--driver-frames", type=int, default=60)
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------------------------
def _gen_frame(job):
    from hive_b200.synthetic import Scene
    config, variant, i, dim = job
    sc = _scene(config, variant, dim)
    fr = sc.frame(i)
    return fr.rgb, fr.depth, fr.pose


def _scene(config, variant, dim=None):
    from hive_b200 import synthetic
    if dim is not None:
        key = f"{config}x{dim[2]}"
        if key not in synthetic.CONFIGS:
            cfg = dict(synthetic.CONFIGS[config])
            cfg["dim"] = tuple(dim)
            synthetic.CONFIGS[key] = cfg
        config = key
    return synthetic.Scene(config, variant)


def make_frames(config, variant, n, dim=None):
    """Generate n synthetic frames on the host cores (process pool; called before CUDA is initialised)."""
    import multiprocessing as mp
    jobs = [(config, variant, i, dim) for i in range(n)]
    procs = min(len(jobs), os.cpu_count() or 1)
    if procs <= 1:
        return [_gen_frame(j) for j in jobs]
    with mp.get_context("fork").Pool(procs) as pool:
        return pool.map(_gen_frame, jobs)


def config_dict(args, sc, dim, n_gpus, n_distinct):
    """`config` of the JSON line -- the same for both arms (what differs, the reference arm's bounded sample, is
    described in its `cpu_baseline.sample`)."""
    return {
        "workload": f"{args.config}: {'x'.join(map(str, dim))} TSDF @ {sc.voxel_size * 100:g} cm, {sc.W}x{sc.H} synthetic "
                    f"frames, scene={args.scene}, {args.batch} frames/step",
        "voxels": int(np.prod(dim)), "frames_per_step": args.batch, "scene": args.scene,
        "parallelism": "single GPU" if n_gpus == 1 else f"z-slab x{n_gpus}, NCCL frame broadcast",
        "l2": "volume planes (12 B/voxel, >= 1.6 GB per GPU) are far larger than the 126 MB L2; no flush needed",
        "distinct_frames": n_distinct,
    }


def global_dim(args, n_gpus):
    from hive_b200 import synthetic
    base_dim = synthetic.CONFIGS[args.config]["dim"]
    if args.scaling == "weak":
        return (base_dim[0], base_dim[1], base_dim[2] * n_gpus)  # one base volume per GPU along z
    return tuple(base_dim)  # BASELINE configs C3 / C5: the named global volume, cut into N z-slabs


def n_distinct_frames(args):
    return max(2, min(args.distinct_frames, args.batch * (args.steps + args.warmup)))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu_index, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "power_w_max": max(power) if power else None, "samples": len(sm)}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic():
    """Per-launch DRAM bytes of the integrate kernel from the committed ncu capture, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "integrate_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


def reference_speed_record():
    """What oracle/measure_reference_speed.py measured in the build container, where /root/reference exists."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_reference_speed.json")) as f:
            rec = json.load(f)
        return {"source": "profiles/r02_reference_speed.json (build container, real reference vs restatements, C1)",
                "cores": rec["cores"], **{k: {m: v[m] for m in ("reference_numpy_numba", "restatement_numba",
                                                                "c_port_openmp", "c_port_vs_reference",
                                                                "numba_restatement_vs_reference")}
                                          for k, v in rec["scenes"].items()}}
    except Exception as e:
        return {"error": str(e)}


def host_cpu():
    try:
        with open("/proc/cpuinfo") as f:
            model = f.read().split("model name")[1].split("\n")[0].strip(": \t")
    except Exception:
        model = "unknown"
    return {"model": model, "logical_cores": os.cpu_count()}


def numba_restatement_speed(variant, frames, n_sample=2):
    """The reference-shaped numpy + numba restatement (oracle/fusion_numba.py) on a bounded sample: `n_sample` frames
    into the central 256^3 voxels of the C2 volume (same 1 cm voxels, same frames)."""
    try:
        import numba
        from oracle import fusion_numba
    except Exception as e:
        return {"numba_imported": False, "error": str(e)}
    sc = _scene("C2", variant)
    lo = sc.vol_bnds[:, 0] + 128 * sc.voxel_size
    bnds = np.stack([lo, lo + (256 - 0.5) * sc.voxel_size], axis=1)
    vol = fusion_numba.NumbaTSDFVolume(bnds, sc.voxel_size)
    rgb, depth, pose = frames[0]
    vol.integrate(rgb, depth, sc.K, pose, 1.0)  # numba JIT + page faults
    t0 = time.perf_counter()
    for i in range(n_sample):
        rgb, depth, pose = frames[(i + 1) % len(frames)]
        vol.integrate(rgb, depth, sc.K, pose, 1.0)
    dt = time.perf_counter() - t0
    return {"numba_imported": True, "numba": numba.__version__, "threads": numba.get_num_threads(),
            "value": float(np.prod(vol._vol_dim)) * n_sample / dt, "unit": UNIT,
            "sample": f"{n_sample} frames of C2/{variant} into the central 256^3 voxels through oracle/fusion_numba.py "
                      f"(the reference's structure: (N,3) voxel table, numba prange kernels, numpy gather/scatter), {dt:.1f} s"}


# ----------------------------------------------------------------------------------------------------------------------
def cpu_baseline(config, variant, frames, n_sample, dim=None):
    """Oracle C port of fusion.py:260-303 on all host cores, bounded sample of the same workload."""
    from oracle import c_oracle
    c_oracle.build()
    c_oracle.use_all_host_threads()
    sc = _scene(config, variant, dim)
    vol = c_oracle.COracleTSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
    n_vox = int(np.prod(vol._vol_dim))
    rgb, depth, pose = frames[0]
    vol.integrate(rgb, depth, sc.K, pose)  # warm-up (page faults of the 3 planes)
    t0 = time.perf_counter()
    for i in range(n_sample):
        rgb, depth, pose = frames[(i + 1) % len(frames)]
        vol.integrate(rgb, depth, sc.K, pose)
    dt = time.perf_counter() - t0
    return {"value": n_vox * n_sample / dt, "unit": UNIT, "cores": c_oracle.num_threads(), "kind": "port",
            "sample": f"{n_sample} frames of {config}/{variant} ({'x'.join(map(str, vol._vol_dim))} voxels, "
                      f"{depth.shape[1]}x{depth.shape[0]}) through oracle/tsdf_oracle.c (OpenMP), {dt:.1f} s",
            "s_per_frame": dt / n_sample, "host": host_cpu(),
            # what a HIVE user without a GPU runs is far slower than this C port: the reference-shaped restatement
            # measured here, and the real reference measured in the build container
            "reference_shaped": numba_restatement_speed(variant, frames),
            "port_vs_reference_speed": reference_speed_record()}


def run_reference(args):
    """The reference arm: the reference's CPU path on the host cores, one frame of the workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import c_oracle
    c_oracle.build()
    c_oracle.use_all_host_threads()
    n_distinct = min(args.distinct_frames, 8)
    frames = make_frames(args.config, args.scene, n_distinct)
    sc = _scene(args.config, args.scene)
    try:
        import numba
        from oracle import fusion_numba
        numba_note = {"numba_imported": True, "numba": numba.__version__, "threads": numba.get_num_threads()}
    except Exception as e:
        fusion_numba, numba_note = None, {"numba_imported": False, "error": str(e)}

    # the C/OpenMP port on the full volume, a few frames: context for the headline below
    cvol = c_oracle.COracleTSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
    n_vox_full = int(np.prod(cvol._vol_dim))
    cvol.integrate(frames[0][0], frames[0][1], sc.K, frames[0][2])
    t0 = time.perf_counter()
    for i in range(4):
        rgb, depth, pose = frames[(i + 1) % len(frames)]
        cvol.integrate(rgb, depth, sc.K, pose)
    c_port = {"value": n_vox_full * 4 / (time.perf_counter() - t0), "unit": UNIT, "cores": c_oracle.num_threads(),
              "sample": f"4 frames into the full {'x'.join(map(str, cvol._vol_dim))} volume through oracle/tsdf_oracle.c"}
    del cvol

    if fusion_numba is not None:
        # bounded sample: the central 256^3 voxels of the volume (the reference's CPU mode needs ~200 B/voxel and
        # ~4 s per frame for them); same voxel size, same frames, one frame per step
        sub = 256 if args.steps + args.warmup <= 40 else 160
        dim0 = np.array(sc.dim)
        sub_dim = np.minimum(dim0, sub)
        lo = sc.vol_bnds[:, 0] + ((dim0 - sub_dim) // 2) * sc.voxel_size
        bnds = np.stack([lo, lo + (sub_dim - 0.5) * sc.voxel_size], axis=1)
        vol = fusion_numba.NumbaTSDFVolume(bnds, sc.voxel_size)
        vol.integrate(frames[0][0], frames[0][1], sc.K, frames[0][2], 1.0)  # numba JIT, excluded like any warm-up
        what = (f"oracle/fusion_numba.py: the reference's CPU path restated with its own structure (fusion.py:166-215, "
                f"260-303: (N,3) voxel table, three numba prange kernels, numpy gather/scatter), "
                f"{numba_note['threads']} numba threads; 0.99-1.06x the real reference's speed where both can run")
        cores = int(numba_note["threads"])
    else:
        vol = c_oracle.COracleTSDFVolume(sc.vol_bnds.copy(), sc.voxel_size)
        what = f"oracle/tsdf_oracle.c, the C/OpenMP port of fusion.py:260-303 (numba does not import on this box)"
        cores = c_oracle.num_threads()
    n_vox = int(np.prod(vol._vol_dim))
    for i in range(args.warmup):
        rgb, depth, pose = frames[i % len(frames)]
        vol.integrate(rgb, depth, sc.K, pose, 1.0)
    t0 = time.perf_counter()
    for i in range(args.steps):
        rgb, depth, pose = frames[(args.warmup + i) % len(frames)]
        vol.integrate(rgb, depth, sc.K, pose, 1.0)
    dt = time.perf_counter() - t0
    value = n_vox * args.steps / dt
    sample = (f"each step = 1 frame (of the GPU arm's {args.batch}-frame step) of {args.config}/{args.scene} into "
              f"{'x'.join(map(str, vol._vol_dim))} voxels (voxel-updates/s is size-normalised) through {what}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 projection / f32 state", "data": "synthetic",
        "config": config_dict(args, sc, global_dim(args, max(args.gpus, 1)), max(args.gpus, 1), n_distinct_frames(args)),
        "frames_per_s": args.steps / dt,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "host": host_cpu(), **numba_note},
        "c_port": c_port,
        "port_vs_reference_speed": reference_speed_record(),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def parity_check(local, sc, frames, sequence, n_blocks=8, block=(16, 16, 64)):
    """Random blocks of the volume the timed loops left behind against the oracle (oracle/tsdf_oracle.c over the blocks'
    slices of the global world tables, the whole frame sequence, obs_weight 1).  Runs after the timing, on rank 0."""
    from oracle import c_oracle
    from hive_b200.fusion import _world_table
    c_oracle.build()
    X, Y, Z = (int(d) for d in local._local_dim)
    bx, by, bz = (min(b, d) for b, d in zip(block, (X, Y, Z)))
    rng = np.random.default_rng(1)
    seen = (local._weight_vol_gpu[::bx, ::by, :Z:bz] > 0).nonzero().cpu().numpy() * np.array([bx, by, bz])
    if len(seen) == 0:
        return {"parity_checked": False, "why": "no voxel was updated"}
    picks = seen[rng.choice(len(seen), min(n_blocks, len(seen)), replace=False)]
    picks = np.minimum(picks, np.array([X - bx, Y - by, Z - bz]))
    tables = [_world_table(local._vol_origin[0], local._voxel_size, X), _world_table(local._vol_origin[1], local._voxel_size, Y),
              _world_table(local._vol_origin[2], local._voxel_size, Z, local._z0)]
    prepared = {i: c_oracle.prepare_frame(frames[i][0], frames[i][1], sc.K, frames[i][2]) for i in sorted(set(sequence))}
    worst, ok_w, ok_c, n_vox, n_upd = 0.0, True, True, 0, 0
    for x0, y0, z0 in picks:
        orc = c_oracle.COracleBlock([tables[0][x0:x0 + bx], tables[1][y0:y0 + by], tables[2][z0:z0 + bz]], local._voxel_size)
        for i in sequence:
            orc.integrate_prepared(prepared[i])
        sl = (slice(x0, x0 + bx), slice(y0, y0 + by), slice(z0, z0 + bz))
        worst = max(worst, float(np.abs(local._tsdf_vol_gpu[sl].cpu().numpy() - orc._tsdf_vol_cpu).max()))
        ok_w &= bool(np.array_equal(local._weight_vol_gpu[sl].cpu().numpy(), orc._weight_vol_cpu))
        ok_c &= bool(np.array_equal(local._color_vol_gpu[sl].cpu().numpy(), orc._color_vol_cpu))
        n_vox += bx * by * bz
        n_upd += int((orc._weight_vol_cpu > 0).sum())
    passed = worst <= 1e-5 and ok_w and ok_c
    return {"parity_checked": bool(passed), "voxels_compared": n_vox, "of_which_updated": n_upd,
            "frames_integrated": len(sequence), "max_abs_tsdf_err": worst, "weight_bit_exact": ok_w,
            "colour_bit_exact": ok_c, "blocks": f"{len(picks)} random {bx}x{by}x{bz} blocks with weight > 0",
            "oracle": "oracle/tsdf_oracle.c (fusion.py:260-303) over the same frame sequence"}


# ----------------------------------------------------------------------------------------------------------------------
def measure_mesh(vol, torch, peak, repeats=5, cpu=True):
    """get_mesh device time, median of `repeats`: `mc_ms` = CUDA events around the five kernels of one extraction (the
    end event is recorded right behind the last launch; capacity known from the warm-up extraction, so nothing on the host
    sits between the kernels); `mc_call_ms` = the same start event to an event recorded after extract_device_mesh
    returned, i.e. including the host's read of the totals (one pinned D2H + stream synchronise)."""
    from hive_b200 import marching_cubes as mc
    times, call_times, mesh = [], [], None
    for r in range(repeats + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        mesh = mc.extract_device_mesh(vol, timing_event=True)
        e1.record()
        torch.cuda.synchronize()
        done = vol._mc_plan.done_event
        if r > 1 and done is not None:
            times.append(e0.elapsed_time(done))
            call_times.append(e0.elapsed_time(e1))
    ms = float(np.median(times))
    V, F = int(mesh.verts.shape[0]), int(mesh.faces.shape[0])
    n = int(np.prod(vol._local_dim))
    algo = 4.0 * n + 4.0 * V + 28.0 * V + 12.0 * F
    e2e = []
    for r in range(4):  # first call sizes the pinned staging blocks; median of the next three
        verts = faces = norms = colors = None  # a caller that drops the previous mesh: its pinned blocks are reused
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        verts, faces, norms, colors = vol.get_mesh()
        e2e.append((time.perf_counter() - t0) * 1000.0)
    e2e_ms = float(np.median(e2e[1:]))
    res = {"mc_ms": ms, "mc_call_ms": float(np.median(call_times)), "vertices": V, "triangles": F, "voxels": n, "algorithmic_bytes": algo,
           "achieved_gbs": algo / (ms / 1e3) / 1e9, "roofline_frac": algo / (ms / 1e3) / 1e9 / peak,
           "get_mesh_e2e_ms": e2e_ms, "d2h_bytes": int(verts.nbytes + faces.nbytes + norms.nbytes + colors.nbytes),
           "scene": "room", "launches_per_extraction": mc.LAUNCHES_PER_EXTRACTION}
    if cpu:
        # the C restatement of the skimage call, single thread, on the same volume -- and the checker
        from oracle import c_oracle, fusion_oracle as fo
        c_oracle.build()
        tsdf, color = vol.get_volume()
        t0 = time.perf_counter()
        ov, of, on, _, ok = c_oracle.marching_cubes(tsdf, 0.0, with_keys=True)
        cpu_s = time.perf_counter() - t0
        order = c_oracle.product_order(ok)
        rank = np.empty_like(order)
        rank[order] = np.arange(len(order))
        ow, oc = fo.mesh_postprocess(ov[order], color, vol._voxel_size, vol._vol_origin)
        same = bool(faces.shape == of.shape and np.array_equal(faces, rank[of].astype(np.int32))
                    and np.array_equal(verts, ow) and np.array_equal(colors, oc))
        res["cpu_baseline"] = {"ms": cpu_s * 1e3, "cores": 1, "kind": "port", "host": host_cpu(),
                               "what": "oracle/mc_oracle.c: serial C restatement of skimage.measure.marching_cubes "
                                       "(fusion.py:338) on the same volume; skimage itself is on neither box",
                               "speedup_device": cpu_s * 1e3 / ms, "speedup_e2e": cpu_s * 1e3 / e2e_ms}
        res["parity_checked"] = same
    return res


def measure_mesh_sharded(vol, world, rank, dist, torch, repeats=3):
    """Seam-placed get_mesh of the z-slab sharded volume: halo exchange + per-slab marching cubes + column-count all-gather
    + rows stored straight into rank 0's HBM (or gathered) + D2H, wall clock between barriers, max over ranks (median of `repeats` after one warm-up)."""
    times, out = [], None
    for r in range(repeats + 1):
        out = None
        _barrier(world, dist, torch)
        t0 = time.perf_counter()
        out = vol.get_mesh()
        torch.cuda.synchronize()
        ms = _max_over_ranks((time.perf_counter() - t0) * 1000.0, world, dist, torch)
        if r > 0:
            times.append(ms)
    res = {"get_mesh_e2e_ms": float(np.median(times)), "scene": "room", "slabs": world,
           "voxels": int(np.prod(vol.global_dim)), "stitched": True,
           # "p2p": every slab's emit kernel stored its rows into rank 0's HBM over NVLink; "gather": grouped send/recv + scatter
           "assembly": getattr(vol, "last_mesh_path", None)}
    vol.close()  # collective: unmap / free the shared mesh buffer
    if rank == 0:
        res.update(vertices=int(out[0].shape[0]), triangles=int(out[1].shape[0]),
                   d2h_bytes=int(sum(a.nbytes for a in out)))
    return res


class _BenchDataset:
    """What hive.fusion.tsdf_fusion reads from a HiveDataset (hive/fusion.py:79-134), over pre-generated host frames."""

    def __init__(self, frames, K, n):
        from hive_b200.synthetic import _Indexed, _Trajectory
        self.num_frames = n
        self.camera_matrix = K
        self.has_inpainted_frame_data = False
        self.camera_trajectory = _Trajectory([np.linalg.inv(frames[i % len(frames)][2]) for i in range(n)])
        H, W = frames[0][1].shape
        v, u = np.mgrid[0:H, 0:W]
        self._masks = [(((u - W * (0.3 + 0.04 * k)) / (0.08 * W)) ** 2 + ((v - 0.55 * H) / (0.3 * H)) ** 2 < 1.0)
                       .astype(np.uint8) for k in range(8)]  # a walking-person sized blob, 8 positions
        self.bg_rgb_dataset = _Indexed(lambda i: frames[i % len(frames)][0])
        self.bg_depth_dataset = _Indexed(lambda i: frames[i % len(frames)][1].copy())
        self.mask_dataset = _Indexed(lambda i: self._masks[i % 8])


def measure_driver(args, base_cfg, frames, torch):
    """The reference's call site end to end: hive.fusion.tsdf_fusion(dataset, options) (hive/fusion.py:79-134) -- frustum
    bounds + voxel budget, per-frame 10x mask dilation + depth masking + integrate from host frames, get_mesh to host."""
    from hive_b200 import hive_fusion, synthetic
    sc = _scene(base_cfg, "room")
    n = args.driver_frames
    ds = _BenchDataset(frames, sc.K, n)
    n_vox = int(np.prod(synthetic.CONFIGS[base_cfg]["dim"]))
    opts = hive_fusion.BackgroundMeshOptions(depth_mask_dilation_iterations=10, sdf_voxel_size=sc.voxel_size,
                                             sdf_max_voxels=n_vox)
    times, mesh = [], None
    for r in range(3):
        mesh = None
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mesh = hive_fusion.tsdf_fusion(ds, opts)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    s = float(np.median(times[1:]))
    vs, bounds = hive_fusion.adjust_voxel_size(ds, opts, range(n))
    dim = np.ceil((bounds[:, 1] - bounds[:, 0]) / vs).astype(int)
    cpu = None
    if not args.no_cpu_baseline:
        # the same per-frame loop on the host cores: cv2.dilate x10 + masking as the reference does it, then the C port
        from oracle import c_oracle, driver_oracle as drv
        c_oracle.build()
        vol = c_oracle.COracleTSDFVolume(bounds.copy(), vs)
        poses = ds.camera_trajectory.inverse().to_homogenous_transforms()
        k = 6
        t0 = time.perf_counter()
        for i in range(k):
            depth = ds.bg_depth_dataset[i]
            depth[drv.dilate_mask(ds.mask_dataset[i], 10) > 0] = 0.0
            vol.integrate(ds.bg_rgb_dataset[i], depth, ds.camera_matrix, poses[i], 1.0)
        cpu = {"frames_per_s": k / (time.perf_counter() - t0), "frames": k, "cores": c_oracle.num_threads(),
               "kind": "port", "what": "per-frame part only (dilate + mask + integrate), no bounds pass, no marching cubes"}
        del vol
    return {"cpu_baseline": cpu, "api": "hive_b200.hive_fusion.tsdf_fusion(dataset, options) == hive.fusion.tsdf_fusion (hive/fusion.py:79-134)",
            "frames": n, "seconds": s, "frames_per_s": n / s, "voxel_updates_per_s": float(np.prod(dim)) * n / s,
            "volume": [int(d) for d in dim], "voxel_size": float(vs), "mask_dilation_iterations": 10,
            "vertices": int(mesh.vertices.shape[0]), "triangles": int(mesh.faces.shape[0]),
            "includes": "frustum bounds over all frames, volume allocation, mask dilation + depth masking (fused into "
                        "the upload), integrate from host frames, marching cubes, mesh D2H"}


def measure_d2m(args, base_cfg, world, rank, torch, dist, peak, repeats=5):
    """Per-frame foreground depth-to-mesh (BASELINE config C4 shape: 640x480, frame-partitioned across ranks)."""
    from hive_b200 import geometric
    B = args.d2m_frames
    sc = _scene(base_cfg, "room")
    n_distinct = min(12, B)
    fr = [sc.frame(i + 1000 * rank) for i in range(n_distinct)]  # every rank owns its own frame range: no collective
    depths = np.stack([fr[i % n_distinct].depth for i in range(B)])
    masks = np.stack([(fr[i % n_distinct].mask > 0) for i in range(B)]).astype(np.uint8)
    poses = np.stack([np.linalg.inv(fr[i % n_distinct].pose) for i in range(B)])
    dev = torch.device("cuda", torch.cuda.current_device())
    d = torch.from_numpy(depths).to(dev)
    m = torch.from_numpy(masks).to(dev)
    cams = torch.from_numpy(np.concatenate([poses[:, :3, :3].reshape(B, 9), poses[:, :3, 3].reshape(B, 3)], axis=1)).to(dev)
    kinv = geometric._kinv64(sc.K)
    times, out = [], None
    for r in range(repeats + 1):
        _barrier(world, dist, torch)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = geometric.depth_to_mesh_device(d, m, kinv, cams, 0.1, 1)
        e1.record()
        torch.cuda.synchronize()
        if r > 0:
            times.append(_max_over_ranks(e0.elapsed_time(e1), world, dist, torch))
    ms = float(np.median(times))
    V, F = int(out.verts.shape[0]), int(out.faces.shape[0])
    H, W = sc.H, sc.W
    algo = 5.0 * H * W * B + 36.0 * V + 12.0 * F  # depth f32 + mask u8 in; f64 xyz + (u, v, depth) + faces out
    # end to end: host numpy in, host numpy out (H2D of depth+mask, D2H of vertices and faces)
    e2e = []
    for r in range(4):  # first call sizes the pinned staging blocks; median of the next three
        hv = hf = hp = hd = None
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        batch = geometric.depth_to_mesh_batch(depths, masks, sc.K, poses)
        hv, hf, hp, hd = batch.to_host()
        e2e.append(time.perf_counter() - t0)
        del batch
    e2e_s = float(np.median(e2e[1:]))
    e2e_batch_fps = _sum_over_ranks(int(B / e2e_s), world, dist, torch)
    # the same through the pipelined host call: chunks of frames, upload / kernels / download overlapped on three streams
    e2e = []
    for r in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        meshes = geometric.depth_to_mesh_host(depths, masks, sc.K, poses)
        e2e.append(time.perf_counter() - t0)
        assert meshes.n_vertices == V and meshes.n_faces == F
        del meshes
    e2e_fps = _sum_over_ranks(int(B / float(np.median(e2e[1:]))), world, dist, torch)
    e2e = []
    for r in range(4):  # and what a glTF export needs: float32 positions + faces only
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        meshes = geometric.depth_to_mesh_host(depths, masks, sc.K, poses, vertex_dtype=np.float32, want_points=False)
        e2e.append(time.perf_counter() - t0)
        del meshes
    e2e_lean_fps = _sum_over_ranks(int(B / float(np.median(e2e[1:]))), world, dist, torch)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the reference's per-frame code (point_cloud_from_depth + scipy Delaunay + _filter_faces,
        # hive/pipeline.py:386-409) single thread on 20 frames, and through a ThreadPool over all cores as
        # hive/pipeline.py:491 runs it
        from multiprocessing.pool import ThreadPool
        from oracle import d2m_oracle as d2m

        def one(i):
            w2c = poses[i]
            return d2m.frame_mesh(depths[i], masks[i] > 0, sc.K, w2c[:3, :3], w2c[:3, 3:])[0].shape[0]

        one(0)
        n1 = min(20, B)
        t0 = time.perf_counter()
        for i in range(n1):
            one(i)
        single = n1 / (time.perf_counter() - t0)
        threads = os.cpu_count() or 1
        t0 = time.perf_counter()
        with ThreadPool(threads) as pool:
            pool.map(one, range(B))
        pooled = B / (time.perf_counter() - t0)
        cpu = {"frames_per_s_single_thread": single, "frames_per_s_threadpool": pooled, "threads": threads,
               "kind": "port", "host": host_cpu(), "frames": [n1, B],
               "what": "oracle/d2m_oracle.py = hive.geometric.point_cloud_from_depth + scipy.spatial.Delaunay + "
                       "Pipeline._filter_faces (hive/pipeline.py:386-409), ThreadPool as hive/pipeline.py:491"}
    pf = None
    if rank == 0 and world == 1:
        # the composed per-frame step (hive/pipeline.py:340-484): objects resident in HBM vs host arrays between the steps
        from hive_b200 import image_processing, pipeline
        dil = image_processing.MaskDilationOptions(num_iterations=2)
        pf = {}
        for name, fn in (("stepwise_ms_per_frame", pipeline.process_frame_stepwise), ("ms_per_frame", pipeline.process_frame)):
            for rep in range(2):  # first pass warms the allocators
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for f in fr[:6]:
                    mesh_f = fn(f.rgb, f.depth, f.mask, np.linalg.inv(f.pose), sc.K, dil)
                pf[name] = 1e3 * (time.perf_counter() - t0) / len(fr[:6])
        pf["vertices_last_frame"] = int(len(mesh_f.vertices))
        pf["api"] = "hive_b200.pipeline.process_frame(rgb, depth, mask_encoded, pose, K, ...): host frame in, FrameMesh (vertices, faces, atlas, uv) out"
    algo_survey = 5.0 * H * W * B + 24.0 * V + 12.0 * F  # without the optional (u, v, depth) outputs
    return {"frames_per_s": world * B / (ms / 1e3), "ms_per_batch": ms, "frames_per_batch_per_gpu": B,
            "vertices": V, "faces": F, "achieved_gbs": algo_survey / (ms / 1e3) / 1e9,
            "roofline_frac": algo_survey / (ms / 1e3) / 1e9 / peak,
            "roofline_frac_with_points2d": algo / (ms / 1e3) / 1e9 / peak,
            "roofline_note": "roofline_frac uses SURVEY 8(d)'s 5*H*W + 24*V + 12*F; the kernel also writes 12 B/vertex of "
                             "(u, v, depth) for _filter_faces' callers, counted in roofline_frac_with_points2d",
            "cpu_baseline": cpu, "e2e_frames_per_s": float(e2e_fps), "e2e_one_batch_frames_per_s": float(e2e_batch_fps),
            "e2e_f32_positions_only_frames_per_s": float(e2e_lean_fps), "d2h_bytes_f32_positions_only": int(12 * V + 12 * F),
            "e2e_api": "hive_b200.geometric.depth_to_mesh_host (host arrays in, per-frame host arrays out, chunk pipeline); "
                       "e2e_one_batch = depth_to_mesh_batch + to_host (upload, kernels, download in sequence)",
            "process_frame": pf,
            "h2d_bytes": int(depths.nbytes + masks.nbytes), "d2h_bytes": int(hv.nbytes + hf.nbytes + hp.nbytes + hd.nbytes),
            "workload": f"{B} frames/GPU of {W}x{H}, instance masks of the room scene, frame-partitioned"}


# ----------------------------------------------------------------------------------------------------------------------
def measure_c5_strong(args, frames, world, rank, torch, dist, Vol, peak):
    """BASELINE config C5 at its named size: ONE 2048^3 volume (103 GB of voxel state), 1920x1080 frames, cut into
    `world` z-slabs (strong scaling).  With more than one GPU the same frames are afterwards integrated into the whole
    volume on rank 0 alone, so that the line carries its own single-GPU baseline (`speedup_vs_n1`)."""
    from hive_b200 import synthetic
    dim = tuple(synthetic.CONFIGS["C5"]["dim"])
    sc = _scene("C5", "dense", dim)
    n_vox = int(np.prod(dim))
    warm, timed = 2, max(2, args.c5_frames)

    sequence = []

    def run(vol, w, r):
        dev_frames = vol.upload_frames(frames if r == 0 else None, sc.K)
        n_f = dev_frames.num_frames
        del sequence[:]
        sequence.extend([i % n_f for i in range(warm)] + [(warm + i) % n_f for i in range(timed)])
        vol.count_updates(False)
        vol.integrate_batch_device(dev_frames, [i % n_f for i in range(warm)])
        vol.count_updates(True)
        vol.reset_update_counter()
        _barrier(w, dist, torch)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(0, timed, 2):
            vol.integrate_batch_device(dev_frames, [(warm + i + j) % n_f for j in range(min(2, timed - i))])
        e1.record()
        torch.cuda.synchronize()
        _barrier(w, dist, torch)
        ms = _max_over_ranks(e0.elapsed_time(e1), w, dist, torch)
        updates = _sum_over_ranks(vol.num_updated_local(), w, dist, torch)
        return ms, updates

    vol = Vol(sc.vol_bnds.copy(), sc.voxel_size, world_size=world, rank=rank)
    ms, updates = run(vol, world, rank)
    value = n_vox * timed / (ms / 1e3)
    per_gpu_gbs = (24.0 * updates / world + 8.0 * sc.H * sc.W * timed) / (ms / 1e3) / 1e9
    res = {"workload": f"C5: {'x'.join(map(str, dim))} TSDF @ {sc.voxel_size * 100:g} cm ({12 * n_vox / 1e9:.0f} GB of voxel state), "
                       f"{sc.W}x{sc.H} frames, scene=dense, {world} z-slab(s)",
           "scaling": "strong", "n_gpus": world, "frames_timed": timed, "value": value, "unit": UNIT,
           "frames_per_s": timed / (ms / 1e3), "ms_per_frame": ms / timed,
           "updated_fraction": updates / float(n_vox * timed), "achieved_gbs_per_gpu": per_gpu_gbs,
           "roofline_frac_per_gpu": per_gpu_gbs / peak}
    if world == 1:
        res["parity"] = parity_check(vol.local, sc, frames, sequence)  # block-sampled against the oracle, incl. z > 1024
    del vol
    torch.cuda.empty_cache()
    if world > 1:
        _barrier(world, dist, torch)
        if rank == 0:
            from hive_b200.sharded import SlabShardedTSDFVolume
            one = SlabShardedTSDFVolume(sc.vol_bnds.copy(), sc.voxel_size, world_size=1, rank=0)
            ms1, _ = run(one, 1, 0)
            res["n1"] = {"value": n_vox * timed / (ms1 / 1e3), "ms_per_frame": ms1 / timed,
                         "what": "the same frames into the whole 2048^3 volume on rank 0 alone, same run"}
            res["speedup_vs_n1"] = res["value"] / res["n1"]["value"]
            res["parity"] = parity_check(one.local, sc, frames, sequence)
            del one
            torch.cuda.empty_cache()
        _barrier(world, dist, torch)
    return res


# ----------------------------------------------------------------------------------------------------------------------
def run_ours(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = max(world, 1)
    base_cfg = args.config

    dim = global_dim(args, n_gpus)
    n_distinct = n_distinct_frames(args)

    # Host-side synthetic data first (fork pool), CUDA afterwards.  Only rank 0 owns frames; others receive broadcasts.
    scenes = [args.scene] + ([] if args.no_room or args.scene == "room" else ["room"])
    frames = {v: make_frames(base_cfg, v, n_distinct, dim if n_gpus > 1 else None) if rank == 0 else None for v in scenes}
    do_c5 = not args.no_c5 and base_cfg == "C2" and args.scaling == "weak"  # the default (driver) invocation only
    from hive_b200 import synthetic as _syn
    c5_frames = make_frames("C5", "dense", 4, tuple(_syn.CONFIGS["C5"]["dim"])) if do_c5 and rank == 0 else None

    import torch
    torch.cuda.set_device(local_rank)
    import torch.distributed as dist
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from hive_b200 import build as hv_build
    if rank == 0:
        hv_build.build()
    if world > 1:
        dist.barrier()
    from hive_b200.sharded import SlabShardedTSDFVolume

    peak, peak_src = measured_peak()
    out = None
    results = {}
    for variant in scenes:
        sc = _scene(base_cfg, variant, dim if n_gpus > 1 else None)
        res = _measure(args, sc, frames[variant], variant, world, rank, local_rank, torch, dist,
                       SlabShardedTSDFVolume, peak)
        results[variant] = res
        torch.cuda.empty_cache()

    d2m_res = None
    if not args.no_d2m:
        d2m_res = measure_d2m(args, base_cfg, world, rank, torch, dist, peak)

    c5_res = None
    if do_c5:
        torch.cuda.empty_cache()
        c5_res = measure_c5_strong(args, c5_frames, world, rank, torch, dist, SlabShardedTSDFVolume, peak)

    driver_res = None
    if not args.no_driver and world == 1 and "room" in frames:
        torch.cuda.empty_cache()
        driver_res = measure_driver(args, base_cfg, frames["room"], torch)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    main = results[args.scene]
    H, W = main["H"], main["W"]
    out = {
        "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f64 projection / f32 state", "data": "synthetic",
        "config": config_dict(args, sc, dim, n_gpus, n_distinct),
        "frames_per_s": main["frames_per_s"],
        "updated_fraction": main["updated_fraction"],
        "per_rank": main.get("per_rank"),
        "e2e": main["e2e"],
        "gpu_launches": main["gpu_launches"],
        "clocks": main["clocks"],
        "roofline": {
            "bound": "hbm", "achieved": main["achieved_gbs"], "peak": peak, "unit": "GB/s",
            "frac": main["achieved_gbs"] / peak, "traffic": None, "peak_source": peak_src,
            "kernel": "hv::integrate_pipe_kernel",
            "algorithmic_bytes_per_launch": main["bytes_per_launch"],
            "avg_launch_ms": main["avg_launch_ms"],
            "frac_of_8000_spec": main["achieved_gbs"] / 8000.0,
        },
    }
    from hive_b200 import _lib as hv_lib
    src_hash = hv_build.source_hash()
    out["build"] = {"binary_id": hv_lib.build_id(), "source_hash": src_hash, "binary_matches_sources": hv_lib.build_id() == src_hash,
                    "library": os.path.relpath(hv_lib.LIB_PATH, ROOT)}
    out["parity_checked"] = bool(main["parity"].get("parity_checked", False))
    out["parity"] = main["parity"]
    tr = load_traffic()
    if tr is not None:
        out["roofline"]["traffic"] = tr.get("dram_bytes_per_launch")
        out["roofline"]["traffic_source"] = tr.get("source")
    if "room" in results and args.scene != "room":
        r = results["room"]
        out["room"] = {k: r[k] for k in ("value", "frames_per_s", "updated_fraction", "achieved_gbs", "avg_launch_ms")}
        out["room"]["roofline_frac"] = r["achieved_gbs"] / peak
        out["room"]["e2e_value"] = r["e2e"]["value"]
    for r in results.values():
        if r.get("mesh") is not None:
            out["mesh"] = r["mesh"]
    if d2m_res is not None:
        out["depth_to_mesh"] = d2m_res
    if c5_res is not None:
        out["c5_strong"] = c5_res
    if driver_res is not None:
        out["driver"] = driver_res
    if not args.no_cpu_baseline and n_gpus == 1:
        out["cpu_baseline"] = cpu_baseline(base_cfg, args.scene, frames[args.scene], args.cpu_sample_frames)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def _measure(args, sc, frames, variant, world, rank, local_rank, torch, dist, Vol, peak):
    B, K, Wm = args.batch, args.steps, args.warmup
    vol = Vol(sc.vol_bnds.copy(), sc.voxel_size, world_size=world, rank=rank)
    n_vox_global = int(np.prod(vol.global_dim))
    H, W = sc.H, sc.W

    # ---- device-resident path -----------------------------------------------------------------------------------
    dev_frames = vol.upload_frames(frames, sc.K)  # rank 0 uploads; other ranks get an empty receiver of the same shape
    n_f = dev_frames.num_frames

    sequence = []  # every frame integrated into this volume, in order (for the parity check after the timing)

    def step_device(s):
        idx = [(s * B + j) % n_f for j in range(B)]
        sequence.extend(idx)
        vol.integrate_batch_device(dev_frames, idx)

    vol.count_updates(False)
    for s in range(Wm):
        step_device(s)
    vol.count_updates(True)
    vol.reset_update_counter()
    _barrier(world, dist, torch)
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(K):
        step_device(Wm + s)
    e1.record()
    torch.cuda.synchronize()
    _barrier(world, dist, torch)
    clocks = sampler.stop() if sampler is not None else None
    ms_local = e0.elapsed_time(e1)
    ms = _max_over_ranks(ms_local, world, dist, torch)
    local_updates = vol.num_updated_local()
    total_updates = _sum_over_ranks(local_updates, world, dist, torch)
    per_rank = None
    if world > 1:  # which slab sets the pace, and how much work each one had (the slabs of one scene are not alike)
        mine = torch.tensor([ms_local, float(local_updates)], dtype=torch.float64, device="cuda")
        every = torch.empty((world, 2), dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(every.view(-1), mine)
        every = every.cpu().numpy()
        per_rank = {"ms_per_step": [float(v) / K for v in every[:, 0]],
                    "updated_fraction": [float(v) / (n_vox_global / world * K * B) for v in every[:, 1]]}
    n_frames = K * B
    value = n_vox_global * n_frames / (ms / 1000.0)

    # roofline of the dominant kernel on this rank: algorithmic bytes / event-timed launch duration
    launches = n_frames
    # (N > 1: the per-GPU average -- slabs of a realistic scene see very different numbers of updates)
    bytes_per_launch = (24.0 * total_updates / world + 8.0 * H * W * n_frames) / launches
    avg_launch_ms = ms / launches
    achieved_gbs = bytes_per_launch / (avg_launch_ms / 1000.0) / 1e9

    # ---- end-to-end path: host numpy frames through TSDFVolume.integrate -------------------------------------------
    vol.count_updates(True)
    for s in range(min(Wm, 3)):
        sequence.extend((s * B + j) % n_f for j in range(B))
        vol.integrate_batch_host(frames, [(s * B + j) % n_f for j in range(B)], sc.K)
    vol.read_update_counter()
    _barrier(world, dist, torch)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e2.record()
    for s in range(K):
        sequence.extend(((Wm + s) * B + j) % n_f for j in range(B))
        vol.integrate_batch_host(frames, [((Wm + s) * B + j) % n_f for j in range(B)], sc.K)
        vol.read_update_counter_pipelined()  # D2H read of the step's result (async snapshot, consumed one step later)
    vol.read_update_counter()  # drains the pipeline: the last step's result
    e3.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1000.0
    _barrier(world, dist, torch)
    ms_e2e = _max_over_ranks(max(e2.elapsed_time(e3), wall_ms), world, dist, torch)
    e2e_value = n_vox_global * n_frames / (ms_e2e / 1000.0)

    parity = {"parity_checked": False, "why": "not rank 0"}
    if rank == 0:
        parity = parity_check(vol.local, sc, frames, sequence)

    mesh = None
    if not args.no_mesh and variant == "room" and world == 1:
        mesh = measure_mesh(vol.local, torch, peak, cpu=not args.no_cpu_baseline)
    elif not args.no_mesh and variant == "room":
        mesh = measure_mesh_sharded(vol, world, rank, dist, torch)

    return {
        "value": value, "ms_per_step": ms / K, "frames_per_s": n_frames / (ms / 1000.0),
        "updated_fraction": total_updates / float(n_vox_global * n_frames),
        "achieved_gbs": achieved_gbs, "bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_launch_ms,
        "gpu_launches": launches, "clocks": clocks, "H": H, "W": W, "mesh": mesh, "parity": parity, "per_rank": per_rank,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * (H * W * 7 + 12 * 8 + 16),
                "d2h_bytes_per_step": 8, "ms_per_step": ms_e2e / K,
                "api": "TSDFVolume.integrate(color_im, depth_im, cam_intr, cam_pose) with host numpy buffers"},
    }


def _barrier(world, dist, torch):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def _max_over_ranks(x, world, dist, torch):
    if world == 1:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _sum_over_ranks(x, world, dist, torch):
    if world == 1:
        return int(x)
    t = torch.tensor([x], dtype=torch.int64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return int(t.item())


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
