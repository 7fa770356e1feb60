"""K1 building block: batched m x m Gauss-Jordan inverse with partial pivoting, singular test and
infinity norm (``inverse_block`` + ``block_norm``, main.cpp:746-820, 669-683)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native


def block_inverse(blocks: torch.Tensor, norm_a: float):
    """``blocks``: (k, m, m) fp64 CUDA tensor.  Returns (inverses, inf-norms, ok-flags)."""
    assert blocks.is_cuda and blocks.dtype == torch.float64 and blocks.dim() == 3
    k, m, m2 = blocks.shape
    assert m == m2
    blocks = blocks.contiguous()
    inv = torch.zeros_like(blocks)
    norms = torch.zeros(k, dtype=torch.float64, device=blocks.device)
    ok = torch.zeros(k, dtype=torch.int32, device=blocks.device)
    lib = _native.cuda()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _native.check(lib.jb_block_inverse(C.c_void_p(blocks.data_ptr()), k, m, C.c_double(norm_a),
                                       C.c_void_p(inv.data_ptr()), C.c_void_p(norms.data_ptr()),
                                       C.c_void_p(ok.data_ptr()), st))
    return inv, norms, ok


def cpu_inverse_block(a: torch.Tensor, norm_a: float):
    """CPU restatement (oracle).  Returns (inverse, singular_flag)."""
    a = a.detach().to(torch.float64).contiguous().clone()
    n = a.shape[0]
    b = torch.empty_like(a)
    sing = _native.host().jb_cpu_inverse_block(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()),
                                               C.c_double(norm_a), n)
    return b, bool(sing)
