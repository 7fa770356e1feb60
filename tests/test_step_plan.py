"""Scheduling invariants of the fused step kernel, checked on the CPU through the same code the
kernel uses (csrc/step_plan.h via jb_plan_enumerate).

Look-ahead schedule per step t:  A(t) -> B1(t) [K1(t+1) beside it] -> B2(t), where B2(t) also
carries SCALE(t+1) on the owner of the next pivot row (the pivot row of step t itself was scaled by
B2(t-1), or by the S prologue for t = 0).

* across A -> B1 -> B2 every (row tile, column tile) of every local block row except row t is
  updated exactly once, and ROWT covers every (sub tile, column tile) once; SCALE(t+1) covers every
  (sub tile, column tile) once, on the next pivot owner only;
* nothing waits for work of a LATER launch, and inside one launch only for work with a LOWER claim
  index (items are claimed in order, so that work is already running):
    - UPD / ROWT of step t need SCALE(t) of their column tile: an earlier launch (B2(t-1) / S);
    - a swapped-in row additionally needs the ROWT pushes of its own column tile and of the tiles
      that hold block column t (this dependency once dead-locked when those pushes lived in a
      later phase);
    - SCALE(t+1) of a column tile needs UPD(t) of the next pivot row on that tile: phase A or B1
      (earlier launches that never wait for B2) or the UPDP items at the head of B2 (lower claim
      index);
* the pivot-row buffer of parity (t+1)&1 that SCALE(t+1) overwrites is no longer read by step t-1
  (its readers are the phases of step t-1, all launched before K1(t+1), whose cross-rank min-loc
  every rank's B2(t) waits for): checked here as "SCALE(t+1) lives in B2(t), never earlier".
"""
import ctypes as C

import pytest

from mpi_jordan_crazy_acceleration_b200 import _native
from mpi_jordan_crazy_acceleration_b200.parallel.layout import num_block_rows, padded_ld, rows_per_process

SCALE, ROWT, UPD, UPD_SW = 0, 1, 2, 3
ALL, A, B, A2, S, B1, B2 = 0, 1, 2, 3, 4, 5, 6
ORDER = {A: 0, B1: 1, B2: 2}
BM, BN = 64, 128


def enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, phase, pct0, pct1, nb1=0,
                    next_rank=-1, next_slot=0, nearly=0):
    cap = 1 << 20
    buf = (C.c_int32 * (5 * cap))()
    n = _native.host().jb_plan_enumerate(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, phase, pct0, pct1,
                                         nb1, nearly, next_rank, next_slot, buf, cap)
    assert n >= 0
    return [tuple(buf[5 * i + k] for k in range(5)) for i in range(n)]


def engine_priority(t, m, CT):
    # mirrors launch_step() in csrc/engine.cu
    return (t * m) // BN, min(CT - 1, ((t + 2) * m - 1) // BN)


CASES = [
    # n, m, p, t, pivot block row of step t (global), pivot block row of step t+1 (global), B1 share
    (1024, 128, 1, 0, 0, 1, 0.4), (1024, 128, 1, 3, 6, 6, 0.4), (1024, 128, 2, 2, 5, 3, 0.25),
    (1536, 128, 3, 4, 4, 9, 0.5), (2048, 256, 2, 1, 6, 6, 0.85), (960, 60, 4, 5, 11, 13, 0.4),
    (512, 32, 8, 3, 9, 4, 0.4), (4096, 128, 8, 10, 25, 11, 0.3), (640, 64, 2, 7, 9, 8, 0.0),
    (300, 7, 3, 20, 40, 21, 1.0), (4096, 128, 8, 10, 10, 26, 0.4), (1024, 128, 2, 2, 2, 5, 0.6),
]


@pytest.mark.parametrize("n,m,p,t,piv_g,nxt_g,frac", CASES)
def test_phases_cover_everything_once_and_never_wait_forward(n, m, p, t, piv_g, nxt_g, frac):
    Nr = num_block_rows(n, m)
    assert t < Nr - 1 and t <= piv_g < Nr and t + 1 <= nxt_g < Nr
    ldw = padded_ld(n, m)
    CT = (ldw + BN - 1) // BN
    nsub = (m + BM - 1) // BM
    piv_rank, piv_slot, do_swap = piv_g % p, piv_g // p, int(piv_g != t)
    nxt_rank, nxt_slot = nxt_g % p, nxt_g // p
    pct0, pct1 = engine_priority(t, m, CT)
    nnp = CT - (pct1 - pct0 + 1)
    nearly = max(0, min(nnp, 8, engine_priority(t + 1, m, CT)[1] - pct1))   # mirrors launch_step()
    nb1 = min(nnp, max(nearly, int(frac * nnp + 0.999)))
    pri_next = set(range(*[(a, b + 1) for a, b in [engine_priority(t + 1, m, CT)]][0]))
    t_rank = t % p
    tcols = set(range((t * m) // BN, ((t + 1) * m - 1) // BN + 1))   # tiles holding block column t

    rowt_at, scale_next_at, updp_at = {}, {}, {}
    per_rank = {}
    for rank in range(p):
        rloc = rows_per_process(Nr, p, rank)
        items = {}
        for ph in (A, B1, B2):
            items[ph] = enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, ph, pct0, pct1,
                                        nb1, nxt_rank, nxt_slot, nearly)
            for idx, (ty, ct, lrow0, rows, prow) in enumerate(items[ph]):
                sub = (lrow0 % m) // BM if ty != SCALE else lrow0 // BM
                if ty == SCALE:
                    assert ph == B2 and rank == nxt_rank        # SCALE(t+1): next owner, never before B2(t)
                    assert (sub, ct) not in scale_next_at
                    scale_next_at[(sub, ct)] = (ORDER[ph], idx)
                elif ty == ROWT:
                    assert rank == t_rank and ph in (A, B1)
                    assert (sub, ct) not in rowt_at
                    rowt_at[(sub, ct)] = (ORDER[ph], idx)
                elif prow:
                    assert ph == B2 and rank == nxt_rank and lrow0 // m == nxt_slot
                    updp_at[(sub, ct)] = (ORDER[ph], idx)
        per_rank[rank] = items
        # same UPD coverage as the un-split kernel
        upd_split = sorted((lrow0, ct) for ph in (A, B1, B2) for ty, ct, lrow0, _, _ in items[ph]
                           if ty in (UPD, UPD_SW))
        upd_all = sorted((lrow0, ct) for ty, ct, lrow0, _, _ in
                         enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, ALL, 0, -1)
                         if ty in (UPD, UPD_SW))
        assert upd_split == upd_all
        assert len(set(upd_split)) == len(upd_split)
        want = {(s * m + sub * BM, ct) for s in range(rloc) for sub in range(nsub) for ct in range(CT)
                if not (rank == t_rank and s == t // p)}
        assert set(upd_split) == want
        # the types agree with the un-split kernel too (swapped-in row)
        ty_split = {(lrow0, ct): ty for ph in (A, B1, B2) for ty, ct, lrow0, _, _ in items[ph] if ty in (UPD, UPD_SW)}
        ty_all = {(lrow0, ct): ty for ty, ct, lrow0, _, _ in
                  enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, ALL, 0, -1)
                  if ty in (UPD, UPD_SW)}
        assert ty_split == ty_all

    full = {(sub, ct) for sub in range(nsub) for ct in range(CT)}
    assert set(rowt_at) == full
    assert set(scale_next_at) == full
    # the tiles phase A of step t+1 needs are scaled first, and none of them waits for the whole of
    # B1: they are priority tiles of step t (finished by A(t)) or B1's flagged early tiles
    first = sorted(scale_next_at.items(), key=lambda kv: kv[1])[:nsub * (pct1 - pct0 + 1 + nearly)]
    assert pri_next <= {ct for (_, ct), _ in first}
    b1_order = []
    for ty, ct, *_ in per_rank[nxt_rank][B1]:
        if ty in (UPD, UPD_SW) and ct not in b1_order:
            b1_order.append(ct)
    for ct in pri_next - set(range(pct0, pct1 + 1)):
        assert ct in b1_order[:nearly], (ct, b1_order[:4], nearly)

    def available(dep, where_rank, me_rank, me):
        """dep = (phase order, idx) of a producer on rank `where_rank`; me likewise on me_rank."""
        if dep[0] != me[0]:
            return dep[0] < me[0]
        return where_rank != me_rank or dep[1] < me[1]   # same launch: lower claim index, or another GPU

    for rank, items in per_rank.items():
        upd_at = {}
        for ph in (A, B1, B2):
            for idx, (ty, ct, lrow0, rows, prow) in enumerate(items[ph]):
                if ty in (UPD, UPD_SW):
                    upd_at[(lrow0, ct)] = (ORDER[ph], idx)
        for ph in (A, B1, B2):
            for idx, (ty, ct, lrow0, rows, prow) in enumerate(items[ph]):
                me = (ORDER[ph], idx)
                sub = (lrow0 % m) // BM
                if ty == UPD_SW:
                    assert do_swap and rank == piv_rank
                    for c in tcols | {ct}:
                        assert available(rowt_at[(sub, c)], t_rank, rank, me), (ph, ct, c)
                if ty == SCALE:
                    # needs UPD(t) of every sub tile of the next pivot row on this column tile
                    for s in range(nsub):
                        dep = upd_at[(nxt_slot * m + s * BM, ct)]
                        assert available(dep, rank, rank, me), (ct, s, dep, me)
                        if dep[0] == ORDER[B2]:      # same launch: must be flagged as a UPDP item
                            assert (s, ct) in updp_at


def test_prologue_and_serial_plans():
    n, m, p = 2048, 128, 2
    Nr, ldw = num_block_rows(n, m), padded_ld(n, m)
    CT = (ldw + BN - 1) // BN
    nsub = 2
    for rank in range(p):
        rloc = rows_per_process(Nr, p, rank)
        s_items = enumerate_items(m, p, rank, rloc, ldw, 0, 1, 3, 1, S, 0, -1)
        if rank == 1:
            assert sorted((lrow0 // BM, ct) for ty, ct, lrow0, _, _ in s_items) == \
                sorted((s, c) for s in range(nsub) for c in range(CT))
            assert all(ty == SCALE for ty, *_ in s_items)
        else:
            assert s_items == []
        tot = lambda ph, a, b, **kw: len(enumerate_items(m, p, rank, rloc, ldw, 3, 1, 2, 1, ph, a, b, **kw))  # noqa: E731
        a, b = engine_priority(3, m, CT)
        # A + B1 + B2 (no next pivot on this rank) == ALL minus its SCALE items
        n_scale = nsub * CT if rank == 1 else 0
        assert tot(A, a, b) + tot(B1, a, b, nb1=5) + tot(B2, a, b, nb1=5) == tot(ALL, 0, -1) - n_scale
        # last step of the look-ahead schedule: one wide phase A = ROWT + UPD of everything
        assert tot(A, 0, CT - 1) == tot(ALL, 0, -1) - n_scale
        # tcgen05 mode: A2 with an empty priority range = SCALE + ROWT of every tile
        a2 = enumerate_items(m, p, rank, rloc, ldw, 3, 1, 2, 1, A2, 0, -1)
        assert sum(1 for ty, *_ in a2 if ty == SCALE) == n_scale
        assert sum(1 for ty, *_ in a2 if ty == ROWT) == (nsub * CT if 3 % p == rank else 0)
        assert all(ty in (SCALE, ROWT) for ty, *_ in a2)
