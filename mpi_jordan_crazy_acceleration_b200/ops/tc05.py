"""K3b building block: ``C += L @ Rn`` on the tcgen05 int8-split path (TMEM accumulators)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native


def tc05_gemm_update(c: torch.Tensor, left: torch.Tensor, right: torch.Tensor, slices: int = 8) -> torch.Tensor:
    """In-place ``c += left @ right`` (fp64 CUDA tensors); ``left.shape[0]`` must be a multiple of 128."""
    assert c.is_cuda and c.dtype == left.dtype == right.dtype == torch.float64
    rows, k = left.shape
    k2, cols = right.shape
    assert k == k2 and c.shape == (rows, cols) and c.stride(1) == 1 and left.stride(1) == 1 and right.stride(1) == 1
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _native.check(_native.cuda().jb_tc05_gemm_update(
        C.c_void_p(c.data_ptr()), c.stride(0), rows, cols, C.c_void_p(left.data_ptr()), left.stride(0),
        C.c_void_p(right.data_ptr()), right.stride(0), k, slices, st))
    return c
