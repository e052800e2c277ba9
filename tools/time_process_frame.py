#!/usr/bin/env python
"""process_frame (objects resident in HBM) against process_frame_stepwise (host arrays between the steps) on C2 room frames."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from hive_b200 import image_processing, pipeline
    from hive_b200.synthetic import Scene
    sc = Scene("C2", "room")
    frames = [sc.frame(i) for i in range(8)]
    dil = image_processing.MaskDilationOptions(num_iterations=2)
    for name, fn in (("stepwise", pipeline.process_frame_stepwise), ("device-resident", pipeline.process_frame)):
        for rep in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            nv = 0
            for fr in frames:
                m = fn(fr.rgb, fr.depth, fr.mask, np.linalg.inv(fr.pose), sc.K, dil)
                nv += len(m.vertices)
            dt = time.perf_counter() - t0
        print(f"{name}: {1e3 * dt / len(frames):.2f} ms per frame ({nv // len(frames)} vertices per frame, "
              f"{len(frames) / dt:.1f} frames/s)", flush=True)


if __name__ == "__main__":
    main()
