"""Scheduling invariants of the fused step kernel, checked on the CPU through the same code the
kernel uses (csrc/step_plan.h via jb_plan_enumerate):

* across the look-ahead phases A1 -> A2 -> B every (row tile, column tile) of every local block row
  except row t is updated exactly once, and SCALE / ROWT cover every (sub tile, column tile) once;
* nothing waits for work of a LATER launch, and inside one launch only for work with a LOWER claim
  index (items are claimed in order, so that work is already running): UPD needs SCALE of its column
  tile; ROWT needs SCALE of its tile; a swapped-in row additionally needs the ROWT pushes of its own
  column tile and of the tiles that hold block column t (this is the dependency that once
  dead-locked when those pushes lived in a later phase).
"""
import ctypes as C

import pytest

from mpi_jordan_crazy_acceleration_b200 import _native
from mpi_jordan_crazy_acceleration_b200.parallel.layout import num_block_rows, padded_ld, rows_per_process

SCALE, ROWT, UPD, UPD_SW = 0, 1, 2, 3
ALL, A1, B, A2 = 0, 1, 2, 3
ORDER = {A1: 0, A2: 1, B: 2}
BM, BN = 64, 128


def enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, phase, pct0, pct1):
    cap = 1 << 20
    buf = (C.c_int32 * (4 * cap))()
    n = _native.host().jb_plan_enumerate(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, phase, pct0, pct1,
                                         buf, cap)
    assert n >= 0
    return [(buf[4 * i], buf[4 * i + 1], buf[4 * i + 2], buf[4 * i + 3]) for i in range(n)]


def engine_priority(t, m, CT):
    # mirrors launch_step() in csrc/engine.cu
    return (t * m) // BN, min(CT - 1, ((t + 2) * m - 1) // BN)


CASES = [
    # n, m, p, t, pivot block row (global)
    (1024, 128, 1, 0, 0), (1024, 128, 1, 3, 6), (1024, 128, 2, 2, 5), (1536, 128, 3, 4, 4),
    (2048, 256, 2, 1, 6), (960, 60, 4, 5, 11), (512, 32, 8, 3, 9), (4096, 128, 8, 10, 25),
    (640, 64, 2, 8, 9), (300, 7, 3, 20, 40),
]


@pytest.mark.parametrize("n,m,p,t,piv_g", CASES)
def test_phases_cover_everything_once_and_never_wait_forward(n, m, p, t, piv_g):
    Nr = num_block_rows(n, m)
    assert t < Nr - 1 and t <= piv_g < Nr
    ldw = padded_ld(n, m)
    CT = (ldw + BN - 1) // BN
    nsub = (m + BM - 1) // BM
    piv_rank, piv_slot, do_swap = piv_g % p, piv_g // p, int(piv_g != t)
    pct0, pct1 = engine_priority(t, m, CT)
    t_rank = t % p
    tcols = set(range((t * m) // BN, ((t + 1) * m - 1) // BN + 1))   # tiles holding block column t

    # where (phase order, claim index) each producer item lives, per rank
    scale_at, rowt_at = {}, {}
    per_rank = {}
    for rank in range(p):
        rloc = rows_per_process(Nr, p, rank)
        items = {}
        for ph in (A1, A2, B):
            items[ph] = enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, ph, pct0, pct1)
            for idx, (ty, ct, lrow0, rows) in enumerate(items[ph]):
                sub = (lrow0 % m) // BM if ty != SCALE else lrow0 // BM
                if ty == SCALE:
                    assert rank == piv_rank
                    assert (sub, ct) not in scale_at
                    scale_at[(sub, ct)] = (ORDER[ph], idx)
                elif ty == ROWT:
                    assert rank == t_rank
                    assert (sub, ct) not in rowt_at
                    rowt_at[(sub, ct)] = (ORDER[ph], idx)
        per_rank[rank] = items
        # same UPD coverage as the un-split kernel
        upd_split = sorted((lrow0, ct) for ph in (A1, A2, B) for ty, ct, lrow0, _ in items[ph] if ty in (UPD, UPD_SW))
        upd_all = sorted((lrow0, ct) for ty, ct, lrow0, _ in
                         enumerate_items(m, p, rank, rloc, ldw, t, piv_rank, piv_slot, do_swap, ALL, 0, -1)
                         if ty in (UPD, UPD_SW))
        assert upd_split == upd_all
        assert len(set(upd_split)) == len(upd_split)
        want = {(s * m + sub * BM, ct) for s in range(rloc) for sub in range(nsub) for ct in range(CT)
                if not (rank == t_rank and s == t // p)}
        assert set(upd_split) == want

    assert set(scale_at) == {(sub, ct) for sub in range(nsub) for ct in range(CT)}
    assert set(rowt_at) == {(sub, ct) for sub in range(nsub) for ct in range(CT)}

    def available(dep, where_rank, me_rank, me):
        """dep = (phase order, idx) of a producer on rank `where_rank`; me likewise on me_rank."""
        if dep[0] != me[0]:
            return dep[0] < me[0]
        return where_rank != me_rank or dep[1] < me[1]   # same launch: lower claim index, or another GPU

    for rank, items in per_rank.items():
        for ph in (A1, A2, B):
            for idx, (ty, ct, lrow0, rows) in enumerate(items[ph]):
                me = (ORDER[ph], idx)
                sub = (lrow0 % m) // BM
                if ty in (UPD, UPD_SW, ROWT):
                    for s in range(nsub):                      # the whole pivot-row tile must have landed
                        assert available(scale_at[(s, ct)], piv_rank, rank, me), (ph, ty, ct)
                if ty == UPD_SW:
                    assert do_swap and rank == piv_rank
                    for c in tcols | {ct}:
                        assert available(rowt_at[(sub, c)], t_rank, rank, me), (ph, ct, c)


def test_plan_matches_unsplit_kernel_item_count():
    n, m, p = 2048, 128, 2
    Nr, ldw = num_block_rows(n, m), padded_ld(n, m)
    CT = (ldw + BN - 1) // BN
    for rank in range(p):
        rloc = rows_per_process(Nr, p, rank)
        tot = lambda ph, a, b: len(enumerate_items(m, p, rank, rloc, ldw, 3, 1, 2, 1, ph, a, b))  # noqa: E731
        a, b = engine_priority(3, m, CT)
        assert tot(A1, a, b) + tot(A2, a, b) + tot(B, a, b) == tot(ALL, 0, -1)
