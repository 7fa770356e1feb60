"""Library-call implementation of the same algorithm: ``torch.distributed`` collectives (NCCL on
GPUs, gloo on CPUs) + ``torch.matmul`` (cuBLAS DGEMM) + batched ``torch.linalg.inv_ex``.

Two roles:
* the in-tree **NCCL + cuBLAS baseline** the fused engine is compared against (BASELINE.json
  north_star: "baseline, not the product");
* a CPU-runnable distributed path (gloo, world_size > 1) that exercises the host-side logic —
  block-cyclic layout, min-loc tie rules, swap bookkeeping — without a GPU.

Collective call sites mirror the reference one to one: pivot min-loc = ``all_gather`` of the
24-byte candidate + rank-ordered fold (C9, ``main.cpp:1074``), pivot-row ``broadcast`` (C10,
``:1097``), block-row swap ``send/recv`` (C11, ``:1120-1130``).  The elimination uses the in-place
2 n^3 form like the engine.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.distributed as dist

from .. import _native
from ..parallel.comm import Comm
from ..parallel.layout import Layout

EPS = 1e-15


def formula_rows(fid: int, n: int, rows: torch.Tensor, device="cpu") -> torch.Tensor:
    """Rows ``rows`` (global indices) of the generator matrix (``f``, main.cpp:47-57)."""
    i = rows.to(torch.float64).to(device)[:, None]
    j = torch.arange(n, dtype=torch.float64, device=device)[None, :]
    if fid == 2:
        return 1.0 / (i + j + 1.0)
    if fid == 3:
        return torch.where(j == (n - 1) - i, i + 1.0, torch.zeros((), dtype=torch.float64, device=device))
    if fid == 4:
        return (torch.minimum(i, j) + 1.0) / (torch.maximum(i, j) + 1.0)
    if fid == 5:
        return torch.where(i == j, torch.full((), float(n), dtype=torch.float64, device=device),
                           1.0 / (1.0 + (i - j).abs()))
    if fid == 6:
        ii = torch.remainder(i + n // 2, n)
        return (ii - j).abs()
    return (i - j).abs()


def minloc_fold(cands: list[tuple[float, int, int]]) -> int:
    """``pivot_op`` (main.cpp:729-735) folded in rank order: returns winning rank or -1.

    ``cands[r] = (norm, nonsing, slot)``.  The higher rank survives unless the lower one is
    non-singular and strictly better, so cross-rank ties go to the highest rank."""
    best = 0
    for r in range(1, len(cands)):
        lo, hi = cands[best], cands[r]
        lo_wins = bool(lo[1]) and (not hi[1] or lo[0] < hi[0])
        if not lo_wins:
            best = r
    return best if cands[best][1] else -1


class EagerJordan:
    def __init__(self, n: int, m: int, comm: Optional[Comm] = None, device: str = "cpu"):
        self.comm = comm or Comm()
        self.layout = Layout(n, m, self.comm.world, self.comm.rank)
        self.device = torch.device(device)
        lay = self.layout
        self.Npad = lay.Nr * lay.m
        self.W = torch.zeros(lay.rows_padded, self.Npad, dtype=torch.float64, device=self.device)
        self.formula = None
        self.pivots: list[int] = []

    # ------------------------------------------------------------ set-up
    def _pad_identity(self) -> None:
        lay = self.layout
        for lr in range(lay.rows_real, lay.rows_padded):
            gi = (lr // lay.m * lay.p + lay.rank) * lay.m + lr % lay.m
            self.W[lr].zero_()
            self.W[lr, gi] = 1.0
        self.W[:, lay.n:] *= 0
        for lr in range(lay.rows_real, lay.rows_padded):
            gi = (lr // lay.m * lay.p + lay.rank) * lay.m + lr % lay.m
            self.W[lr, gi] = 1.0

    def fill_formula(self, fid: int) -> None:
        lay = self.layout
        self.formula = fid
        rows = torch.tensor(lay.global_rows(), dtype=torch.int64)
        self.W.zero_()
        if lay.rows_real:
            self.W[: lay.rows_real, : lay.n] = formula_rows(fid, lay.n, rows, self.device)
        self._pad_identity()
        self.A0 = self.W.clone()
        self._norm()

    def upload_local_rows(self, rows: torch.Tensor) -> None:
        lay = self.layout
        self.W.zero_()
        if lay.rows_real:
            self.W[: lay.rows_real, : lay.n] = rows.to(self.device)
        self._pad_identity()
        self.A0 = self.W.clone()
        self._norm()

    def _norm(self) -> None:
        lay = self.layout
        loc = float(self.W[: lay.rows_real, : lay.n].abs().sum(dim=1).max()) if lay.rows_real else 0.0
        self.norm_a = self.comm.allreduce_max(loc)

    # ------------------------------------------------------------ one candidate inversion
    def _invert_blocks(self, blocks: torch.Tensor):
        """Returns (inverses, norms, ok) for a batch of m x m blocks."""
        k, m, _ = blocks.shape
        if self.device.type == "cpu":
            lib = _native.host()
            inv = torch.empty_like(blocks)
            ok = torch.zeros(k, dtype=torch.bool)
            for b in range(k):
                a = blocks[b].contiguous().clone()
                out = torch.empty(m, m, dtype=torch.float64)
                sing = lib.jb_cpu_inverse_block(C.c_void_p(a.data_ptr()), C.c_void_p(out.data_ptr()),
                                                C.c_double(self.norm_a), m)
                ok[b] = sing == 0
                inv[b] = out
            norms = inv.abs().sum(dim=2).max(dim=1).values
            return inv, norms, ok
        inv, info = torch.linalg.inv_ex(blocks)
        ok = (info == 0) & torch.isfinite(inv).all(dim=(1, 2))
        # reference-style relative threshold on the pivots is approximated by the LU diagonal
        lu = torch.linalg.lu_factor_ex(blocks)[0]
        dmin = torch.diagonal(lu, dim1=1, dim2=2).abs().min(dim=1).values
        ok &= dmin >= EPS * self.norm_a
        norms = inv.abs().sum(dim=2).max(dim=1).values
        return inv, norms, ok

    # ------------------------------------------------------------ the elimination
    def invert(self) -> None:
        lay, comm = self.layout, self.comm
        m, p, k, Nr = lay.m, lay.p, lay.rank, lay.Nr
        W = self.W
        self.pivots = []
        use_dist = comm.enabled and comm.world > 1
        for t in range(Nr):
            s0 = max(0, -(-(t - k) // p))
            ncand = max(0, lay.rloc - s0)
            cand = (0.0, 0, s0)
            inv = None
            if ncand:
                blocks = W[s0 * m:, t * m:(t + 1) * m].reshape(ncand, m, m)
                inv, norms, ok = self._invert_blocks(blocks)
                best = -1
                for b in range(ncand):  # strict <, first wins (main.cpp:1051-1064)
                    if bool(ok[b]) and (best < 0 or float(norms[b]) < float(norms[best])):
                        best = b
                if best >= 0:
                    cand = (float(norms[best]), 1, s0 + best)
            cands = comm.allgather_obj(cand) if use_dist else [cand]
            win = minloc_fold(cands)
            if win < 0:
                raise _native.SingularMatrixError(_native.JB_ERR_SINGULAR, "singular matrix")
            piv_slot = cands[win][2]
            piv_g = piv_slot * p + win
            self.pivots.append(piv_g)
            k_t, t_slot = t % p, t // p
            # scaled pivot row, column slot t <- P^-1
            R = torch.empty(m, self.Npad, dtype=torch.float64, device=self.device)
            if k == win:
                Pinv = inv[piv_slot - s0]
                R.copy_(Pinv @ W[piv_slot * m:(piv_slot + 1) * m])
                R[:, t * m:(t + 1) * m] = Pinv
            if use_dist:
                dist.broadcast(R, src=win)
            # swap block rows t <-> piv
            if piv_g != t:
                if k_t == win:
                    if k == win:
                        W[piv_slot * m:(piv_slot + 1) * m] = W[t_slot * m:(t_slot + 1) * m]
                elif k == k_t:
                    dist.send(W[t_slot * m:(t_slot + 1) * m].contiguous(), dst=win)
                elif k == win:
                    buf = torch.empty(m, self.Npad, dtype=torch.float64, device=self.device)
                    dist.recv(buf, src=k_t)
                    W[piv_slot * m:(piv_slot + 1) * m] = buf
            if k == k_t:
                W[t_slot * m:(t_slot + 1) * m] = 0.0
            L = W[:, t * m:(t + 1) * m].clone()
            W[:, t * m:(t + 1) * m] = 0.0
            W.addmm_(L, R, alpha=-1.0)
            if k == k_t:
                W[t_slot * m:(t_slot + 1) * m] = R
        # undo the row interchanges on the columns, last swap first
        for t in range(Nr - 1, -1, -1):
            pg = self.pivots[t]
            if pg != t:
                a = W[:, t * m:(t + 1) * m].clone()
                W[:, t * m:(t + 1) * m] = W[:, pg * m:(pg + 1) * m]
                W[:, pg * m:(pg + 1) * m] = a

    # ------------------------------------------------------------ results
    def local_inverse(self) -> torch.Tensor:
        lay = self.layout
        return self.W[: lay.rows_real, : lay.n]

    def gather_inverse(self) -> Optional[torch.Tensor]:
        lay = self.layout
        parts = self.comm.allgather_obj((lay.global_rows(), self.local_inverse().cpu().numpy()))
        if self.comm.rank != 0:
            return None
        out = torch.empty(lay.n, lay.n, dtype=torch.float64)
        for rows, data in parts:
            if len(rows):
                out[torch.tensor(rows)] = torch.from_numpy(data)
        return out

    def residual(self) -> float:
        """Ring-free residual: every rank gathers X (small n on CPU, baseline on GPU)."""
        lay = self.layout
        parts = self.comm.allgather_obj((lay.global_rows(), self.local_inverse().cpu().numpy()))
        X = torch.empty(lay.n, lay.n, dtype=torch.float64)
        for rows, data in parts:
            if len(rows):
                X[torch.tensor(rows)] = torch.from_numpy(data)
        A = self.A0[: lay.rows_real, : lay.n].cpu()
        D = A @ X
        for i, gi in enumerate(lay.global_rows()):
            D[i, gi] -= 1.0
        loc = float(D.abs().sum(dim=1).max()) if lay.rows_real else 0.0
        return self.comm.allreduce_max(loc)
