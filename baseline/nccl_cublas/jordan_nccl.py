"""In-tree NCCL + cuBLAS baseline: the SAME algorithm as the engine (block Gauss-Jordan with block-row
pivoting, in-place 2n^3 form, 1-D block-cyclic rows), written the way one would with libraries only:

  step t  (reference Jordan(), main.cpp:1026-1194)
    candidates   one batched LU-based inversion of all local candidate blocks     cuSOLVER/cuBLAS batched
    min-loc      all_gather of a 2-double device tensor, arg-min on the device     NCCL   (main.cpp:1074)
    pivot row    owner: P^-1 * row (one GEMM), broadcast                            cuBLAS + NCCL (main.cpp:1097)
    swap         owner of row t sends it to the pivot owner                         NCCL send/recv (main.cpp:1120-1130)
    update       W -= L * row  as ONE addmm_ over the whole local slab             cuBLAS DGEMM (main.cpp:1165-1194)

No pickles, no Python loops over candidates, no per-element host reads: the only host round trip per
step is the 3-int pivot decision (which rank broadcasts), as in the reference, where the result of
MPI_Allreduce lives on the host.  It is the BASELINE the hand-written engine is measured against
(BASELINE.json north star), not a product path.  Works on gloo/CPU tensors too (tests).
"""
from __future__ import annotations

import os
import sys
from typing import Optional

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from mpi_jordan_crazy_acceleration_b200.models.eager import formula_rows  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.parallel.layout import Layout  # noqa: E402

EPS = 1e-15  # main.cpp:7


class SingularMatrix(RuntimeError):
    pass


class NcclCublasJordan:
    def __init__(self, n: int, m: int, device: str = "cuda", group=None):
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        m = min(m, n)
        self.lay = Layout(n, m, self.world, self.rank)
        self.n, self.m, self.dev = n, m, torch.device(device)
        self.Nr = self.lay.Nr
        self.np_ = self.Nr * m                      # padded order (identity padding like the engine)
        self.rloc = self.lay.rloc
        self.W = torch.zeros(self.rloc * m, self.np_, dtype=torch.float64, device=self.dev)
        self.rowbuf = torch.empty(m, self.np_, dtype=torch.float64, device=self.dev)
        self.swapbuf = torch.empty(m, self.np_, dtype=torch.float64, device=self.dev)
        self.pivots: list[int] = []
        self.norm_a = 0.0

    # ------------------------------------------------------------------ set-up
    def global_rows_padded(self) -> torch.Tensor:
        s = torch.arange(self.rloc, dtype=torch.int64)[:, None]
        r = torch.arange(self.m, dtype=torch.int64)[None, :]
        return ((s * self.world + self.rank) * self.m + r).reshape(-1)

    def fill_formula(self, fid: int) -> None:
        gi = self.global_rows_padded()
        real = gi < self.n
        self.W.zero_()
        if real.any():
            rows = formula_rows(fid, self.n, gi[real], device=self.dev)
            self.W[real.to(self.dev).nonzero().squeeze(1), : self.n] = rows
        pad = (~real).nonzero().squeeze(1)
        if pad.numel():
            self.W[pad.to(self.dev), gi[pad].to(self.dev)] = 1.0
        loc = self.W[:, : self.n].abs().sum(1).max() if self.rloc else torch.zeros((), dtype=torch.float64, device=self.dev)
        loc = loc.reshape(1).clone()
        if self.world > 1:
            dist.all_reduce(loc, op=dist.ReduceOp.MAX, group=self.group)
        self.norm_a = float(loc.item())

    # ------------------------------------------------------------------ the elimination
    def invert(self) -> None:
        m, p, k, W = self.m, self.world, self.rank, self.W
        self.pivots = []
        big = torch.tensor(float("inf"), dtype=torch.float64, device=self.dev)
        gather = [torch.empty(2, dtype=torch.float64, device=self.dev) for _ in range(p)]
        for t in range(self.Nr):
            c0, c1 = t * m, (t + 1) * m
            s0 = (t - k + p - 1) // p if t > k else 0          # first local slot with global index >= t
            ncand = self.rloc - s0
            mine = torch.stack([big, torch.zeros((), dtype=torch.float64, device=self.dev)])
            inv = None
            if ncand > 0:
                cand = W[s0 * m:, c0:c1].reshape(ncand, m, m)
                inv, info = torch.linalg.inv_ex(cand)                      # ONE batched factorisation
                norms = inv.abs().sum(2).amax(1)
                # singular test of inverse_block (main.cpp:782): a pivot below EPS * ||A|| shows up as
                # ||inv|| above 1 / (EPS * ||A||); LU break-downs come back through `info`
                bad = (info != 0) | ~torch.isfinite(norms) | (norms * (EPS * self.norm_a) > 1.0)
                norms = torch.where(bad, big, norms)
                best = torch.argmin(norms)                                  # first minimum: lowest slot
                mine = torch.stack([norms[best], (best + s0).to(torch.float64)])
            if p > 1:
                dist.all_gather(gather, mine, group=self.group)            # C9: 16 bytes per rank
                allc = torch.stack(gather)
            else:
                allc = mine[None]
            # rank-ordered fold of pivot_op: strictly smaller wins, ties -> highest rank
            vals = allc[:, 0]
            vmin = vals.min()
            piv_rank_t = (vals == vmin).nonzero().max()
            decision = torch.stack([piv_rank_t.to(torch.float64), allc[piv_rank_t, 1], vmin]).cpu()  # host trip
            if not torch.isfinite(decision[2]):
                raise SingularMatrix(f"singular matrix (step {t})")
            piv_rank, piv_slot = int(decision[0]), int(decision[1])
            piv_g = piv_slot * p + piv_rank
            self.pivots.append(piv_g)
            t_rank, t_slot = t % p, t // p
            # ---- pivot row: scale on the owner, broadcast (C10)
            if k == piv_rank:
                pinv = inv[piv_slot - s0]
                prow = W[piv_slot * m:(piv_slot + 1) * m]
                torch.matmul(pinv, prow, out=self.rowbuf)
                self.rowbuf[:, c0:c1] = pinv                                 # in-place form: slot t <- P^-1
            if p > 1:
                dist.broadcast(self.rowbuf, src=piv_rank, group=self.group)
            # ---- swap: old block row t goes to the pivot row's slot (C11)
            if piv_g != t:
                if t_rank == piv_rank:
                    if k == t_rank:
                        W[piv_slot * m:(piv_slot + 1) * m] = W[t_slot * m:(t_slot + 1) * m]
                else:
                    if k == t_rank:
                        dist.send(W[t_slot * m:(t_slot + 1) * m].contiguous(), dst=piv_rank, group=self.group)
                    if k == piv_rank:
                        dist.recv(self.swapbuf, src=t_rank, group=self.group)
                        W[piv_slot * m:(piv_slot + 1) * m] = self.swapbuf
            # ---- trailing update of every local row except row t: ONE rank-m DGEMM
            if self.rloc:
                L = W[:, c0:c1].clone()
                W[:, c0:c1] = 0.0                                            # slot t starts from B's zero column
                if k == t_rank:
                    L[t_slot * m:(t_slot + 1) * m] = 0.0
                W.addmm_(L, self.rowbuf, alpha=-1.0)
                if k == t_rank:
                    W[t_slot * m:(t_slot + 1) * m] = self.rowbuf
        # column un-permutation of the in-place form (reverse swap order)
        for t in range(self.Nr - 1, -1, -1):
            g = self.pivots[t]
            if g != t and self.rloc:
                a = W[:, t * m:(t + 1) * m].clone()
                W[:, t * m:(t + 1) * m] = W[:, g * m:(g + 1) * m]
                W[:, g * m:(g + 1) * m] = a

    # ------------------------------------------------------------------ results
    def local_inverse(self) -> torch.Tensor:
        gi = self.global_rows_padded()
        return self.W[(gi < self.n).to(self.dev).nonzero().squeeze(1), : self.n]

    def gather_inverse(self) -> Optional[torch.Tensor]:
        gi = self.global_rows_padded()
        loc = self.local_inverse().cpu()
        rows = gi[gi < self.n]
        if self.world == 1:
            out = torch.empty(self.n, self.n, dtype=torch.float64)
            out[rows] = loc
            return out
        parts: list = [None] * self.world
        dist.all_gather_object(parts, (rows, loc), group=self.group)
        if self.rank != 0:
            return None
        out = torch.empty(self.n, self.n, dtype=torch.float64)
        for r, x in parts:
            if len(r):
                out[r] = x
        return out
