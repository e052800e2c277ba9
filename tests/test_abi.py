"""The C-ABI library builds, loads and exports every symbol include/hive_b200.h declares (no compute: no GPU here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "hive_b200.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\bint\s+(hv_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = _declared_symbols()
    for required in ("hv_version", "hv_last_error", "hv_integrate", "hv_pack_frame", "hv_upload_frame"):
        assert required in syms


def test_library_exports_every_declared_symbol(hv_lib):
    for name in _declared_symbols():
        assert hasattr(hv_lib, name), f"{name} declared in include/hive_b200.h but not exported"


def test_python_binding_covers_the_header(hv_lib):
    from hive_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_version_and_error_text(hv_lib):
    assert hv_lib.hv_version() >= 100
    buf = ctypes.create_string_buffer(64)
    assert hv_lib.hv_last_error(buf, 64) >= 0


def test_invalid_arguments_are_rejected_without_a_gpu(hv_lib):
    """Argument validation happens before any CUDA call, so it can be exercised on the CPU-only box."""
    from hive_b200 import _lib
    import pytest
    rc = hv_lib.hv_pack_frame(None, None, 4, 4, None, None)
    assert rc == _lib.HV_E_INVALID
    with pytest.raises(AssertionError):
        _lib.check(rc)
    assert "null" in _lib.last_error()


def test_sources_are_sm100a_only():
    from hive_b200 import build
    flags = " ".join(build.NVCC_FLAGS)
    assert "arch=compute_100a,code=sm_100a" in flags and "-lineinfo" in flags


def test_product_does_not_import_the_oracle():
    """The product path must never route through oracle/ (it is the checker, not a fallback)."""
    pkg = os.path.join(ROOT, "hive_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn
                assert "libhive_oracle" not in src, fn
