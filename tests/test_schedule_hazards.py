"""Race detection without a GPU: the happens-before model of the look-ahead schedule
(tools/schedule_model.py) must find every pair of conflicting accesses ordered - W tiles, pivot-row
buffers on every rank, L panels, candidate inverses, stepinfo - for the shipped schedule and its
switches, at several rank counts.  The work items come from the kernel's own work-list code.

The checker has teeth: with either of the two fixes of this round disabled it reports exactly the
races that had shown up on the GPUs as rare wrong inverses."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from schedule_model import ScheduleModel  # noqa: E402


@pytest.mark.parametrize("Nr,p", [(6, 1), (8, 2), (9, 3), (12, 4), (16, 8), (5, 8)])
@pytest.mark.parametrize("kw", [dict(), dict(sched=1), dict(k1_gated=True), dict(frac=0.03), dict(frac=0.85),
                                dict(sched=1, b2_overlap=True), dict(b2_overlap=True, mma_waits_early=True)])
def test_no_unordered_conflicting_accesses(Nr, p, kw):
    mdl = ScheduleModel(Nr, p, **kw)
    assert mdl.hazards() == []


def test_model_finds_the_missing_step_retired_handshake():
    """Phase A and K1 run ahead of the bulk: without the step-retired tags the owner of pivot(t+2) may
    overwrite a pivot-row buffer that a slower rank still reads for step t."""
    hz = ScheduleModel(8, 3, stepdone_handshake=False).hazards()
    assert hz and all("rowbuf" in h and "SCALE" in h for h in hz)


def test_model_finds_the_uncounted_row_install():
    """When B2 / phase A may start while B1 is still running (JB_B2_OVERLAP=1), B1's early-tile flag
    must also count the install of block row t: phase A of the next step updates that row and K1
    copies it into the L panel.  (Behind a finished B1 - the default - the question does not arise.)"""
    assert ScheduleModel(8, 3, count_rowt_in_early=False).hazards() == []
    hz = ScheduleModel(8, 3, count_rowt_in_early=False, b2_overlap=True, mma_waits_early=True).hazards()
    assert hz and all("ROWT" in h for h in hz)


@pytest.mark.parametrize("Nr,p,seed", [(8, 2, 1), (9, 3, 2), (12, 4, 3), (16, 8, 4), (12, 4, 5), (10, 1, 6)])
@pytest.mark.parametrize("kw", [dict(), dict(sched=1)])
def test_block_row_swaps_are_ordered_too(Nr, p, seed, kw):
    """Random pivot sequences off the diagonal: pushes of the old row t into the pivot owner's swap
    buffer (peer stores + swap flags), update of the swapped-in row from that buffer, next pivot row
    that is itself a swapped-in row - all conflicting accesses ordered."""
    import random

    rnd = random.Random(seed)
    piv = [rnd.choice([t, rnd.randrange(t, Nr)]) for t in range(Nr)]
    assert ScheduleModel(Nr, p, pivots=piv, **kw).hazards() == []


def test_model_finds_the_unguarded_swap_push():
    """Two swaps of the same buffer parity onto one pivot owner: without waiting for that owner's
    step-retired tag the second push may overwrite rows the owner still reads for the first swap."""
    piv = [4, 1, 8, 3, 4, 5, 6, 7, 8, 9, 10, 11]
    assert ScheduleModel(12, 4, pivots=piv).hazards() == []
    hz = ScheduleModel(12, 4, pivots=piv, swap_handshake=False).hazards()
    assert hz and all("swapbuf" in h and "push" in h for h in hz)


def test_model_explains_the_withdrawn_overlap_switch():
    """JB_B2_OVERLAP=1 gave wrong inverses at 8 GPUs (profiles/README.md).  Why: in phase A the flag of
    B1's early tile is awaited by the TMA producer thread only, while the MMA warps read the C tile
    from W as soon as the item is handed to them - unordered with B1's update of that tile once B1
    may still be running.  Behind a finished B1 (the default) and in the in-line schedule there is no
    such pair; with the MMA warps waiting too the overlap would be clean."""
    hz = ScheduleModel(16, 8, b2_overlap=True).hazards()
    assert hz and all("C-read" in h and " A(" in h and " B1(" in h for h in hz)
    assert ScheduleModel(16, 8, b2_overlap=True, mma_waits_early=True).hazards() == []
    assert ScheduleModel(16, 8).hazards() == []


@pytest.mark.parametrize("Nr,p,m", [(12, 2, 60), (12, 3, 64), (8, 2, 256), (9, 4, 256), (16, 8, 32), (10, 3, 96)])
def test_other_block_sizes_straddling_the_column_tiles(Nr, p, m):
    """m != 128: block columns straddle the 128-wide column tiles (m = 60 is BASELINE config 1, m = 256
    config 3, m = 32 config 5), priority ranges and early tiles change shape from step to step."""
    import random

    rnd = random.Random(Nr * 100 + m)
    piv = [rnd.choice([t, rnd.randrange(t, Nr)]) for t in range(Nr)]
    assert ScheduleModel(Nr, p, m=m).hazards() == []
    assert ScheduleModel(Nr, p, m=m, pivots=piv).hazards() == []
    assert ScheduleModel(Nr, p, m=m, pivots=piv, sched=1).hazards() == []
