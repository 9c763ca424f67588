"""CPU checks of the drop-in boundary (no compute calls, no GPU): libgsx.so loads, exports every symbol include/gsx.h
declares and every reference-named C++ function include/gsplat_cuda/*.cuh declares, the ctypes binding lists exactly the
header's functions, host-only queries answer, and the product path fails loudly when the library is missing."""
import ctypes as C
import glob
import importlib
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ffi = importlib.import_module("3dgs_b200._ffi")


def _header_functions():
    src = open(os.path.join(ROOT, "include", "gsx.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)                      # comments mention functions too
    return sorted(set(re.findall(r"\b(gsx_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_function_of_gsx_h():
    L = ffi.lib()
    names = _header_functions()
    assert len(names) >= 50
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, f"declared in include/gsx.h but not exported by libgsx.so: {missing}"


def test_ctypes_binding_lists_exactly_the_header():
    assert sorted(ffi.SYMBOLS) == _header_functions()


def test_library_exports_the_reference_named_cpp_functions():
    """include/gsplat_cuda/*.cuh: the reference's own function names (C++ linkage), defined by compat.cu"""
    declared = set()
    for h in glob.glob(os.path.join(ROOT, "include", "gsplat_cuda", "*.cuh")):
        src = re.sub(r"//.*", "", open(h).read())
        declared |= set(re.findall(r"^(?:void|float)\s+([a-z_0-9]+)\s*\(", src, flags=re.M))
    assert {"render_image", "render_image_backward", "get_sorted_gaussian_list", "adam_step", "fused_loss",
            "compute_morton_codes", "clone_gaussians", "split_gaussians"} <= declared and len(declared) == 20
    out = subprocess.run(["nm", "-DC", "--defined-only", ffi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T ([a-z_0-9]+)\(", out))
    assert declared <= exported, f"declared but not defined: {sorted(declared - exported)}"


def test_host_only_queries_answer_without_a_gpu():
    L = ffi.lib()
    assert L.gsx_version() >= 1
    assert isinstance(L.gsx_last_error(), bytes)
    assert L.gsx_fused_loss_workspace_bytes(16, 16) == 256 + 9 * 16 * 16 * 4
    assert L.gsx_sort_workspace_bytes(C.c_longlong(1_000_000)) > 0
    assert L.gsx_morton_order_workspace_bytes(1000) >= 1000 * (8 + 8 + 4)
    assert L.gsx_sorted_list_workspace_bytes(1000, C.c_longlong(8000)) > 0


def test_argument_errors_are_status_codes_not_crashes():
    L = ffi.lib()
    rc = L.gsx_adam_step_fused(None, 3, None, 10, C.c_float(0.9), C.c_float(0.999), C.c_float(1e-8), C.c_float(0.1),
                               C.c_float(0.001), None, None, None, None)
    assert rc == 1 and b"null" in L.gsx_last_error()                      # GSX_E_NULL
    rc = L.gsx_adam_step_fused(None, 99, None, 10, C.c_float(0.9), C.c_float(0.999), C.c_float(1e-8), C.c_float(0.1),
                               C.c_float(0.001), None, None, None, None)
    assert rc in (1, 3)


def test_missing_library_is_an_import_error_not_a_fallback(monkeypatch):
    monkeypatch.setattr(ffi, "_lib", None)
    monkeypatch.setattr(ffi, "LIB_PATH", os.path.join(ROOT, "3dgs_b200", "no_such_libgsx.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        ffi.lib()
