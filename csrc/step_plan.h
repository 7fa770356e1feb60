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
};

struct Plan {
  int nscale, nrowt, nupd, total;
  int rtu;   // row tiles taking part in UPD
  int npri;  // priority column tiles (they cover block column t+1, needed first by K1(t+1))
};

// Work list per phase (items are claimed IN ORDER from a global counter, so anything an item waits
// for inside the same launch has a lower index and is already running):
//   ALL : SCALE, ROWT, UPD over every column tile                        (no look-ahead)
//   A   : SCALE + ROWT + UPD of the PRIORITY column tiles only (block column t+1): the minimum that
//         K1(t+1) needs, so the pivot search of the next step can start after a ~30 us kernel
//   A2  : SCALE + ROWT of all other column tiles (owner of the pivot row / of row t only)
//   B   : UPD of all other column tiles; runs right behind A2 and is flag-synchronised with it tile
//         by tile, also across GPUs, while K1(t+1) runs beside it on another stream
JB_HD Plan make_plan(const StepParams& P, const JbStepInfo& si) {
  Plan pl;
  pl.npri = (P.phase == JB_PHASE_ALL || P.pct1 < P.pct0) ? 0 : (P.pct1 - P.pct0 + 1);
  pl.rtu = (P.rloc - (P.t_slot >= 0 ? 1 : 0)) * P.nsub;
  const bool owner = (si.piv_rank == P.rank);
  const bool has_t = (P.t_slot >= 0);
  int cts = P.CT;  // column tiles this phase scales / installs
  if (P.phase == JB_PHASE_A) cts = pl.npri;
  else if (P.phase == JB_PHASE_A2) cts = P.CT - pl.npri;
  if (P.phase == JB_PHASE_B) {
    pl.nscale = 0;
    pl.nrowt = 0;
    pl.nupd = pl.rtu * (P.CT - pl.npri);
  } else {
    pl.nscale = owner ? P.nsub * cts : 0;
    pl.nrowt = has_t ? P.nsub * cts : 0;
    pl.nupd = (P.phase == JB_PHASE_A2) ? 0 : pl.rtu * (P.phase == JB_PHASE_ALL ? P.CT : pl.npri);
  }
  pl.total = pl.nscale + pl.nrowt + pl.nupd;
  return pl;
}

// position -> column tile for the SCALE / ROWT items of this phase
JB_HD int ct_of_pos(const StepParams& P, const Plan& pl, int pos) {
  if (P.phase == JB_PHASE_A) return P.pct0 + pos;
  if (P.phase == JB_PHASE_A2) return pos < P.pct0 ? pos : pos + pl.npri;
  return pos;
}

JB_HD Item decode(const StepParams& P, const JbStepInfo& si, const Plan& pl,
                                       int idx) {
  Item it;
  if (idx < pl.nscale) {
    it.type = IT_SCALE;
    const int pos = idx / P.nsub;
    it.sub = idx - pos * P.nsub;
    it.ct = ct_of_pos(P, pl, pos);
    it.lrow0 = it.sub * JB_BM;  // row inside the pivot-row buffer
  } else if (idx < pl.nscale + pl.nrowt) {
    it.type = IT_ROWT;
    const int pos = (idx - pl.nscale) / P.nsub;
    it.sub = (idx - pl.nscale) - pos * P.nsub;
    it.ct = ct_of_pos(P, pl, pos);
    it.lrow0 = P.t_slot * P.m + it.sub * JB_BM;
  } else {
    int u = idx - pl.nscale - pl.nrowt;
    const int cpos = u / pl.rtu;
    int rt = u - cpos * pl.rtu;
    if (P.phase == JB_PHASE_A) it.ct = P.pct0 + cpos;
    else if (P.phase == JB_PHASE_B) it.ct = cpos < P.pct0 ? cpos : cpos + pl.npri;
    else it.ct = cpos;
    int slot = rt / P.nsub;
    it.sub = rt - slot * P.nsub;
    if (P.t_slot >= 0 && slot >= P.t_slot) slot += 1;  // skip block row t (handled by ROWT)
    it.lrow0 = slot * P.m + it.sub * JB_BM;
    it.type = (si.do_swap && si.piv_rank == P.rank && slot == si.piv_slot) ? IT_UPD_SWAPPED : IT_UPD;
  }
  it.rows_valid = ((P.m - it.sub * JB_BM) < JB_BM ? (P.m - it.sub * JB_BM) : JB_BM);
  return it;
}

}  // namespace jbplan
