"""B200-native instant-NGP NeRF hot path: CUDA kernels behind the C ABI of include/ngp_b200.h, the ctypes binding
(`capi`) and the pybind11 `pyngp` module that mirrors the reference's Testbed surface."""
from . import capi  # noqa: F401
