// Work-list logic of the fused step kernel (which items exist in which phase, in which order),
// shared by the device code (csrc/step_fused.cu) and a host-side enumerator used by the CPU tests
// (tests/test_step_plan.py): the scheduling invariants - every tile updated exactly once per step,
// nothing waits for an item of a LATER launch - are checked without a GPU.
#pragma once
#include "engine_types.h"
#include "kernels.h"

#ifdef __CUDACC__
#define JB_HD __host__ __device__ __forceinline__
#else
#define JB_HD inline
#endif

namespace jbplan {

enum { IT_SCALE = 0, IT_ROWT = 1, IT_UPD = 2, IT_UPD_SWAPPED = 3 };

struct Item {
  int type;
  int ct;          // column tile
  int sub;         // 64-row sub tile inside the block row
  int lrow0;       // first local row of the tile in W (output rows / C rows)
  int rows_valid;  // <= JB_BM
  int prow;        // UPD item of the NEXT pivot row (phase B2): publishes prowflag when stored
  int epos;        // >= 0: UPD item of B1's early column tile number epos (counted when stored)
};

struct Plan {
  int nupdp, nscale, nrowt, nupd, nscale2, total;
  int nupd_a;  // B2: bulk UPD items claimed before the second SCALE group
  int rtu;     // row tiles taking part in UPD of step t on this rank
  int rtu2;    // row tiles of the bulk UPD list of this phase (B2 on the next owner: rtu - nsub)
  int npri;    // priority column tiles (they cover block columns t and t+1)
  int nnp;     // non-priority column tiles
  int c0, c1;  // non-priority positions [c0, c1) whose UPD items belong to this phase
  int skip_a, skip_b;  // block-row slots left out of the bulk UPD list, ascending, -1 = none
};

// Work list per phase.  Items are claimed IN ORDER from a global counter, so anything an item waits
// for inside the same launch has a lower index and is already running.
//
// Serial schedule (no look-ahead):
//   ALL : SCALE(t), ROWT(t), UPD(t) over every column tile
// Look-ahead schedule, per step t (the pivot row of step t was scaled and broadcast EARLIER, by
// phase B2 of step t-1, or by the S prologue for t = 0):
//   A   : ROWT(t) + UPD(t) of the PRIORITY column tiles: the minimum K1(t+1) needs, so the pivot
//         search of the next step starts after a short kernel
//   B1  : ROWT(t) of the other column tiles, UPD(t) of the first nb1 of them; K1(t+1) runs beside
//         it on a high-priority stream.  The non-priority tiles are visited cyclically starting
//         right behind the priority range, so B1 begins with the tiles of block column t+2 - the
//         ones phase A of the NEXT step needs; B1 flags each of these `nearly` tiles when done.
//   B2  : launched once K1(t+1) has chosen the next pivot row and B1 has finished (or, with the
//         debugging switch JB_B2_OVERLAP, beside the tail of B1: hence the flags below).  On the
//         owner of that row: UPD(t) of the row's tiles of this phase first (UPDP); then SCALE(t+1)
//         straight into all ranks' pivot-row buffers of step t+1 - the broadcast of step t+1
//         overlaps the bulk of step t - in this order: priority tiles, B1's early tiles (flagged),
//         this phase's tiles; then, on every rank, the bulk of UPD(t), with the SCALE(t+1) items
//         of B1's remaining tiles slipped in after the first few bulk items (B1 has retired by
//         then, so no CTA sits waiting for another launch).
// tcgen05 mode keeps A2 (SCALE(t) + ROWT(t) of every tile) and B (fp64 update of the swapped-in
// row only).
JB_HD Plan make_plan(const StepParams& P, const JbStepInfo& si, const JbStepInfo& sn) {
  Plan pl;
  const int ph = P.phase;
  pl.npri = (ph == JB_PHASE_ALL || ph == JB_PHASE_S || P.pct1 < P.pct0) ? 0 : (P.pct1 - P.pct0 + 1);
  pl.nnp = P.CT - pl.npri;
  pl.rtu = (P.rloc - (P.t_slot >= 0 ? 1 : 0)) * P.nsub;
  pl.rtu2 = pl.rtu;
  pl.skip_a = P.t_slot;
  pl.skip_b = -1;
  const bool owner = (si.piv_rank == P.rank);
  const bool has_t = (P.t_slot >= 0);
  pl.nupdp = pl.nscale = pl.nrowt = pl.nscale2 = 0;
  pl.nupd_a = 0;
  pl.c0 = 0;
  pl.c1 = 0;
  int ucols = 0;  // column tiles of the bulk UPD list
  if (ph == JB_PHASE_ALL) {
    pl.nscale = owner ? P.nsub * P.CT : 0;
    pl.nrowt = has_t ? P.nsub * P.CT : 0;
    pl.c1 = pl.nnp;
    ucols = pl.nnp;
  } else if (ph == JB_PHASE_S) {
    pl.nscale = owner ? P.nsub * P.CT : 0;
  } else if (ph == JB_PHASE_A) {
    pl.nrowt = has_t ? P.nsub * pl.npri : 0;
    ucols = pl.npri;
  } else if (ph == JB_PHASE_A2) {
    pl.nscale = owner ? P.nsub * pl.nnp : 0;
    pl.nrowt = has_t ? P.nsub * pl.nnp : 0;
  } else if (ph == JB_PHASE_B) {
    pl.c1 = pl.nnp;
    ucols = pl.nnp;
  } else if (ph == JB_PHASE_B1) {
    pl.nrowt = has_t ? P.nsub * pl.nnp : 0;
    pl.c1 = P.nb1 < pl.nnp ? P.nb1 : pl.nnp;
    ucols = pl.c1;
  } else {  // JB_PHASE_B2
    pl.c0 = P.nb1 < pl.nnp ? P.nb1 : pl.nnp;
    pl.c1 = pl.nnp;
    ucols = pl.c1 - pl.c0;
    if (P.has_next && sn.piv_rank == P.rank) {
      const int ne = P.nearly < pl.c0 ? P.nearly : pl.c0;
      pl.nupdp = P.nsub * ucols;
      pl.nscale = P.nsub * (pl.npri + ne + ucols);
      pl.nscale2 = P.nsub * (pl.c0 - ne);
      pl.rtu2 = pl.rtu - P.nsub;
      // two slots are left out of the bulk list: block row t (ROWT) and the next pivot row (UPDP)
      if (pl.skip_a < 0 || sn.piv_slot < pl.skip_a) {
        pl.skip_b = pl.skip_a;
        pl.skip_a = sn.piv_slot;
      } else {
        pl.skip_b = sn.piv_slot;
      }
      if (pl.skip_a < 0) {
        pl.skip_a = pl.skip_b;
        pl.skip_b = -1;
      }
    }
  }
  pl.nupd = pl.rtu2 * ucols;
  if (pl.nscale2 > 0) {  // about 1/8 of the bulk, whole column tiles
    const int cols_a = (ucols + 7) / 8;
    pl.nupd_a = pl.rtu2 * cols_a;
  }
  pl.total = pl.nupdp + pl.nscale + pl.nrowt + pl.nupd + pl.nscale2;
  return pl;
}

// position among the non-priority column tiles -> column tile: cyclic, starting right behind the
// priority range (block column t+2 first)
JB_HD int np_ct(const StepParams& P, const Plan& pl, int pos) {
  if (pl.npri == 0) return pos;
  const int ct = P.pct1 + 1 + pos;
  return ct < P.CT ? ct : ct - P.CT;
}
// ... and back (ct must not be a priority tile)
JB_HD int np_pos(const StepParams& P, const Plan& pl, int ct) {
  if (pl.npri == 0) return ct;
  return ct > P.pct1 ? ct - P.pct1 - 1 : ct + P.CT - P.pct1 - 1;
}

JB_HD int upd_type(const StepParams& P, const JbStepInfo& si, int slot) {
  return (si.do_swap && si.piv_rank == P.rank && slot == si.piv_slot) ? IT_UPD_SWAPPED : IT_UPD;
}

JB_HD Item decode(const StepParams& P, const JbStepInfo& si, const JbStepInfo& sn, const Plan& pl,
                  int idx) {
  Item it;
  it.prow = 0;
  it.epos = -1;
  const int ph = P.phase;
  if (idx < pl.nupdp) {  // B2: the next pivot row first
    const int pos = idx / P.nsub;
    it.sub = idx - pos * P.nsub;
    it.ct = np_ct(P, pl, pl.c0 + pos);
    it.lrow0 = sn.piv_slot * P.m + it.sub * JB_BM;
    it.type = upd_type(P, si, sn.piv_slot);
    it.prow = 1;
  } else if (idx < pl.nupdp + pl.nscale) {
    it.type = IT_SCALE;
    const int k = idx - pl.nupdp;
    const int pos = k / P.nsub;
    it.sub = k - pos * P.nsub;
    if (ph == JB_PHASE_B2) {
      // priority tiles, then B1's early tiles (both needed by phase A of the next step), then ours
      const int ne = P.nearly < pl.c0 ? P.nearly : pl.c0;
      if (pos < pl.npri) it.ct = P.pct0 + pos;
      else if (pos < pl.npri + ne) it.ct = np_ct(P, pl, pos - pl.npri);
      else it.ct = np_ct(P, pl, pl.c0 + pos - pl.npri - ne);
    } else {
      it.ct = (ph == JB_PHASE_A2) ? np_ct(P, pl, pos) : pos;
    }
    it.lrow0 = it.sub * JB_BM;  // row inside the pivot-row buffer
  } else if (pl.nscale2 > 0 && idx >= pl.nupdp + pl.nscale + pl.nupd_a &&
             idx < pl.nupdp + pl.nscale + pl.nupd_a + pl.nscale2) {
    // B2: SCALE(t+1) of the tiles phase B1 updates (behind its early tiles)
    it.type = IT_SCALE;
    const int k = idx - (pl.nupdp + pl.nscale + pl.nupd_a);
    const int pos = k / P.nsub;
    it.sub = k - pos * P.nsub;
    const int ne = P.nearly < pl.c0 ? P.nearly : pl.c0;
    it.ct = np_ct(P, pl, ne + pos);
    it.lrow0 = it.sub * JB_BM;
  } else if (idx < pl.nupdp + pl.nscale + pl.nrowt) {
    it.type = IT_ROWT;
    const int k = idx - pl.nupdp - pl.nscale;
    const int pos = k / P.nsub;
    it.sub = k - pos * P.nsub;
    it.ct = (ph == JB_PHASE_A) ? P.pct0 + pos : np_ct(P, pl, pos);
    it.lrow0 = P.t_slot * P.m + it.sub * JB_BM;
  } else {
    int u = idx - pl.nupdp - pl.nscale - pl.nrowt;
    if (pl.nscale2 > 0 && u >= pl.nupd_a) u -= pl.nscale2;  // B2: second part of the bulk
    const int cpos = u / pl.rtu2;
    const int rt = u - cpos * pl.rtu2;
    it.ct = (ph == JB_PHASE_A) ? P.pct0 + cpos : np_ct(P, pl, pl.c0 + cpos);
    int slot = rt / P.nsub;
    it.sub = rt - slot * P.nsub;
    if (pl.skip_a >= 0 && slot >= pl.skip_a) slot += 1;  // block row t (ROWT) / next pivot row (UPDP)
    if (pl.skip_b >= 0 && slot >= pl.skip_b) slot += 1;
    it.lrow0 = slot * P.m + it.sub * JB_BM;
    it.type = upd_type(P, si, slot);
    if (ph == JB_PHASE_B1 && cpos < P.nearly) it.epos = cpos;
  }
  it.rows_valid = ((P.m - it.sub * JB_BM) < JB_BM ? (P.m - it.sub * JB_BM) : JB_BM);
  return it;
}

}  // namespace jbplan
