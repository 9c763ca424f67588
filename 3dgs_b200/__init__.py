"""3dgs_b200 -- B200-native differentiable Gaussian-splatting hot path behind the reference's gsplat_cuda API.

Import with importlib (the package name starts with a digit):  gs = importlib.import_module("3dgs_b200").
Sub-modules: _ffi (ctypes binding of libgsx.so), host (torch-tensor operator interface + Rasterizer), dp (view-parallel
multi-GPU step), scenes (synthetic scenes / cameras).
"""
from . import _ffi, scenes  # noqa: F401

__all__ = ["_ffi", "scenes", "host", "dp"]


def __getattr__(name):
    if name in ("host", "dp"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
