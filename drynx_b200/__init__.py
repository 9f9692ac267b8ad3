"""drynx_b200: B200-native (sm_100a) backend for Drynx's ElGamal + pairing range-proof hot path.

The product is `libdrynx_gpu.so` (CUDA kernels behind the C ABI of include/drynx_gpu.h); this package is the
thin Python host mirror used by tests and bench.py. Importing it never touches oracle/."""
from ._lib import DrynxGpuError, load  # noqa: F401
from .backend import Backend  # noqa: F401
