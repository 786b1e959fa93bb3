"""vt_b200 -- B200-native drop-in for vt's decompose/normalize hot path (C ABI in include/vtn.h,
CUDA kernels in vt_b200/csrc).  The Python layer is plumbing for tests and benchmarks."""
from .api import Batch, Context, NormResult, VtnError  # noqa: F401
