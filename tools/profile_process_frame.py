#!/usr/bin/env python
"""cProfile of hive_b200.pipeline.process_frame on C2 room frames: where the host time of one frame goes."""
import cProfile
import os
import pstats
import sys

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
    args = [(fr.rgb, fr.depth, fr.mask, np.linalg.inv(fr.pose), sc.K, dil) for fr in frames]
    for a in args:
        pipeline.process_frame(*a)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(5):
        for a in args:
            pipeline.process_frame(*a)
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats("cumulative").print_stats(45)


if __name__ == "__main__":
    main()
