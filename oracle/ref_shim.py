"""Import the *real* reference modules from /root/reference (TEST INFRASTRUCTURE, build container only).

``/root/reference`` does not exist on the GPU box, so nothing here may be used by ``-m gpu`` tests, ``smoke()``
or ``bench.py``.  It is used by ``oracle/make_golden.py`` to dump golden vectors and by the CPU-side tests that
pin the restatements in this directory against the reference itself (skipped when the tree is absent).

Shims: ``skimage`` is not installed -> register empty modules before importing
``third_party/tsdf_fusion_python/fusion.py`` (its line 7 is the only obstacle); ``hive.geometric`` needs a
``matplotlib.pyplot`` stub and the numpy-1 aliases ``np.alltrue`` / ``np.product``.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("HIVE_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "third_party", "tsdf_fusion_python", "fusion.py"))


def reference_fusion():
    """The reference's ``fusion`` module (CPU mode only: PyCUDA is absent, fusion.py:9-18)."""
    if "skimage" not in sys.modules:
        sk = types.ModuleType("skimage")
        sk.measure = types.ModuleType("skimage.measure")
        sys.modules["skimage"] = sk
        sys.modules["skimage.measure"] = sk.measure
    path = os.path.join(REFERENCE_ROOT, "third_party", "tsdf_fusion_python")
    if path not in sys.path:
        sys.path.insert(0, path)
    import logging
    lvl = logging.getLogger().level
    logging.getLogger().setLevel(logging.ERROR)
    try:
        import fusion  # noqa: the reference module
    finally:
        logging.getLogger().setLevel(lvl)
    return fusion


def reference_geometric():
    """The reference's ``hive.geometric`` module."""
    import numpy as np
    if not hasattr(np, "alltrue"):
        np.alltrue = np.all
    if not hasattr(np, "product"):
        np.product = np.prod
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        plt.Axes = object
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import hive.geometric as geometric
    return geometric


def reference_pipeline_statics(names=("_triangulate_faces", "_filter_faces", "_get_mesh_texture_and_uv", "_pack_textures")):
    """The reference's own ``Pipeline`` static methods, executed from ``hive/pipeline.py`` where it lies.

    ``hive.pipeline`` cannot be imported here (trimesh, openmesh, detectron2 ...), but these methods only use numpy,
    scipy's Delaunay and ``hive.geometric`` / ``hive.utils`` helpers: their ``FunctionDef`` nodes are taken out of the
    parsed source (decorators dropped) and compiled in a namespace holding exactly those names.  Nothing is copied into
    this repository; the functions exist only in the calling process."""
    import ast
    import numpy as np
    from scipy.spatial import Delaunay
    geometric = reference_geometric()
    import hive.utils as utils  # noqa: importable once hive.geometric is (same shims)
    path = os.path.join(REFERENCE_ROOT, "hive", "pipeline.py")
    with open(path) as f:
        tree = ast.parse(f.read(), filename=path)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "Pipeline")
    picked = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(picked) == len(names), f"missing in {path}: {set(names) - {n.name for n in picked}}"
    for n in picked:
        n.decorator_list = []
    module = ast.Module(body=picked, type_ignores=[])
    ns = {"np": np, "Delaunay": Delaunay, "world2image": geometric.world2image,
          "validate_shape": utils.validate_shape, "validate_camera_parameter_shapes": utils.validate_camera_parameter_shapes,
          "MeshFilteringOptions": object}
    exec(compile(module, path, "exec"), ns)
    return {name: ns[name] for name in names}


def _compile_from_source(rel_path, names, namespace, class_name=None):
    """Compile the named functions (or methods of ``class_name``) of a reference source file in ``namespace``."""
    import ast
    path = os.path.join(REFERENCE_ROOT, rel_path)
    with open(path) as f:
        tree = ast.parse(f.read(), filename=path)
    body = tree.body
    if class_name is not None:
        body = next(n for n in body if isinstance(n, ast.ClassDef) and n.name == class_name).body
    picked = [n for n in body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(picked) == len(names), f"missing in {path}: {set(names) - {n.name for n in picked}}"
    for n in picked:
        n.decorator_list = []
    exec(compile(ast.Module(body=picked, type_ignores=[]), path, "exec"), namespace)
    return {name: namespace[name] for name in names}


def reference_tsdf_driver(volume_class=None):
    """The reference's own ``adjust_voxel_size`` and ``tsdf_fusion`` (hive/fusion.py:37-134), compiled from the source
    file (the module imports trimesh and hive.io at the top and cannot be imported) over the real reference
    ``fusion`` module (CPU mode), the real ``hive.image_processing.dilate_mask`` and ``hive.options``.

    ``volume_class``: replaces ``fusion.TSDFVolume`` inside the driver (a subclass of the reference's, used by the golden
    generator to keep hold of the volume and to answer ``get_mesh``, which needs the absent scikit-image)."""
    import logging
    import typing
    import numpy as np
    fusion = reference_fusion()
    reference_geometric()  # numpy-1 aliases + matplotlib stub, puts the reference root on sys.path
    import hive.image_processing as image_processing
    import hive.options as options
    fusion_ns = types.SimpleNamespace(TSDFVolume=volume_class or fusion.TSDFVolume, get_view_frustum=fusion.get_view_frustum)
    trimesh = types.SimpleNamespace(Trimesh=lambda **kw: kw)
    ns = {"np": np, "logging": logging, "tqdm": lambda it, *a, **k: it, "fusion": fusion_ns, "trimesh": trimesh,
          "dilate_mask": image_processing.dilate_mask, "MaskDilationOptions": options.MaskDilationOptions,
          "BackgroundMeshOptions": options.BackgroundMeshOptions, "HiveDataset": object, "Optional": typing.Optional,
          "List": typing.List, "Tuple": typing.Tuple}
    out = _compile_from_source(os.path.join("hive", "fusion.py"), ("adjust_voxel_size", "tsdf_fusion"), ns)
    out["BackgroundMeshOptions"] = options.BackgroundMeshOptions
    out["fusion"] = fusion
    return out


def reference_select_key_frames():
    """``HiveDataset.select_key_frames`` (hive/io.py:1117-1189) as a plain function of a stand-in ``self``."""
    import logging
    import typing
    import numpy as np
    geo = reference_geometric()
    ns = {"np": np, "logging": logging, "tqdm": lambda it, *a, **k: it, "List": typing.List,
          "get_pose_components": geo.get_pose_components, "pose_vec2mat": geo.pose_vec2mat,
          "point_cloud_from_depth": geo.point_cloud_from_depth, "world2image": geo.world2image}
    return _compile_from_source(os.path.join("hive", "io.py"), ("select_key_frames",), ns,
                                class_name="HiveDataset")["select_key_frames"]
