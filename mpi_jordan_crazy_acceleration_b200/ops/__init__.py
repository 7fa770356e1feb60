"""Thin Python wrappers over individual native kernels (tests, tools, notebooks)."""
from .block_inverse import block_inverse, cpu_inverse_block  # noqa: F401
from .oracle import cpu_block_jordan, cpu_minloc_fold, cpu_residual  # noqa: F401
from .tc05 import tc05_gemm_update  # noqa: F401
from .ozaki_model import split_matmul  # noqa: F401
