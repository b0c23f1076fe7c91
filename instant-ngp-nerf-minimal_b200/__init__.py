"""B200-native instant-NGP NeRF hot path: CUDA kernels behind the C ABI of include/ngp_b200.h, the ctypes binding
(`capi`) and the pybind11 `pyngp` module that mirrors the reference's Testbed surface."""
from . import capi  # noqa: F401


def load_pyngp():
    """the pybind11 `pyngp` module (host/python_api.cpp); raises if it has not been built -- there is no Python fallback"""
    import importlib
    return importlib.import_module(__name__ + ".pyngp")
