"""CPU-side check of the drop-in boundary: the shared library loads without a GPU and exports every
symbol include/psana_ray_cuda.h declares, and the ctypes table mirrors the header.  No compute calls."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "psana_ray_cuda.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(psr_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    import sys
    sys.path.insert(0, os.path.join(ROOT, "psana-ray_b200"))
    import build as psr_build
    psr_build.build_library()
    from psana_ray import _native
    return _native.load()


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("psr_create", "psr_destroy", "psr_set_constants", "psr_set_common_mode", "psr_set_geometry",
                 "psr_calib_batch", "psr_calib_batch_host", "psr_ipc_export_mem", "psr_ipc_open_mem", "psr_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"


def test_ctypes_table_matches_header(lib):
    from psana_ray import _native
    assert sorted(_native.SIGNATURES) == declared_symbols()
    assert lib.psr_version() == 1


def test_missing_library_fails_loudly(tmp_path):
    from psana_ray import _native
    with pytest.raises(_native.NativeLibraryError):
        _native.load(str(tmp_path / "nope.so"))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "psana-ray_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), encoding="utf-8", errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle/"
