"""CPU oracle bindings (``csrc/host/cpu_oracle.cpp``)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native


def cpu_block_jordan(a: torch.Tensor, m: int, p: int = 1):
    """Intended-semantics block Gauss-Jordan emulating ``p`` ranks.  Returns (X, pivots)."""
    n = a.shape[0]
    work = a.detach().to(torch.float64).contiguous().clone()
    x = torch.empty_like(work)
    Nr = (n + m - 1) // m
    piv = (C.c_int32 * Nr)()
    rc = _native.host().jb_cpu_block_jordan(C.c_void_p(work.data_ptr()), C.c_void_p(x.data_ptr()), n, m, p, piv)
    if rc == _native.JB_ERR_SINGULAR:
        raise _native.SingularMatrixError(rc, "singular matrix")
    return x, list(piv)


def cpu_minloc_fold(norms, nonsing) -> int:
    p = len(norms)
    a = (C.c_double * p)(*norms)
    b = (C.c_int32 * p)(*[int(x) for x in nonsing])
    return int(_native.host().jb_cpu_minloc_fold(a, b, p))


def cpu_residual(a: torch.Tensor, x: torch.Tensor) -> float:
    a = a.detach().to(torch.float64).contiguous()
    x = x.detach().to(torch.float64).contiguous()
    return float(_native.host().jb_cpu_residual(C.c_void_p(a.data_ptr()), C.c_void_p(x.data_ptr()), a.shape[0]))
