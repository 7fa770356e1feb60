"""CPU model of the error-free int8 split used by the tcgen05 update path (``csrc/tc05_update.cu``).

The same arithmetic as ``k_slice_rows`` / ``k_slice_cols`` / the epilogue, written with torch integer
tensors: rows of ``L`` and columns of ``R`` are scaled by a power of two, cut into 7-bit signed
digits ``q_s`` with ``|q_s| <= 64`` (radix 128, round to nearest), all slice products are summed
EXACTLY in integers per diagonal ``d = s + u``, and the diagonals are recombined in fp64:

    L R = 2^(eL + eR - 12) * sum_d 2^(-7 d) * sum_{s+u=d} qL_s qR_u

Used by the CPU tests to pin down the algorithm (digit ranges, int32 head-room, accuracy vs slice
count) independently of any GPU.
"""
from __future__ import annotations

import torch


def _exponent(amax: torch.Tensor) -> torch.Tensor:
    """e with amax = f * 2^e, f in [0.5, 1) (0 for amax == 0)."""
    _, e = torch.frexp(amax)
    return torch.where(amax == 0, torch.zeros_like(e), e)


def slice_rows(mat: torch.Tensor, slices: int):
    """Returns (digits [S, rows, K] int64, scale [rows] = 2^(e-6))."""
    e = _exponent(mat.abs().amax(dim=1))
    t = torch.ldexp(mat, (6 - e)[:, None])
    digits = []
    for _ in range(slices):
        q = torch.round(t)          # |q| <= 64; ties-to-even like rint()
        digits.append(q.to(torch.int64))
        t = (t - q) * 128.0         # exact
    return torch.stack(digits), torch.ldexp(torch.ones_like(e, dtype=torch.float64), e - 6)


def split_matmul(left: torch.Tensor, right: torch.Tensor, slices: int = 8):
    """``left @ right`` through the integer split.  Returns (product, max |diagonal accumulator|)."""
    ql, sl = slice_rows(left, slices)
    qr, sr = slice_rows(right.t().contiguous(), slices)    # column scaling of `right`
    acc_max = 0
    out = torch.zeros(left.shape[0], right.shape[1], dtype=torch.float64)
    for d in range(slices - 1, -1, -1):                    # Horner from the smallest diagonal
        dd = torch.zeros(left.shape[0], right.shape[1], dtype=torch.int64)
        for s in range(d + 1):
            dd += ql[s] @ qr[d - s].t()                    # exact integer accumulation
        acc_max = max(acc_max, int(dd.abs().max()))
        out = out * 2.0 ** -7 + dd.to(torch.float64)
    return out * sl[:, None] * sr[None, :], acc_max
