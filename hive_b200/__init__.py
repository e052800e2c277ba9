"""hive_b200 -- B200-native (sm_100a) implementation of HIVE's geometric hot path.

TSDF integration, marching-cubes mesh extraction and per-frame depth-to-mesh behind the reference's own Python API
(``hive_b200.fusion`` mirrors ``third_party/tsdf_fusion_python/fusion.py``; ``hive_b200.geometric`` and
``hive_b200.pipeline`` mirror the hot functions of ``hive/geometric.py`` and ``hive/pipeline.py``).
Python host -> C ABI (``include/hive_b200.h``, ``libhive_b200.so``) -> hand-written CUDA kernels.  No CPU fallback.
"""
__version__ = "0.1.0"
