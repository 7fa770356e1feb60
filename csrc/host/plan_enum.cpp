// Host-side enumeration of the fused step kernel's work list (see csrc/step_plan.h).
#include <string.h>
#include "../step_plan.h"
#include "../../include/jordan_b200.h"

extern "C" {

// Fills out[5*i .. 5*i+4] = {type, column tile, first local row, rows, next-pivot-row flag} for
// every item of the phase, in claim order.  Returns the item count (or -1 if `cap` is too small).
// nb1 / next_piv_*: look-ahead phases B1 / B2 (next_piv_rank < 0: step t is the last one).
int64_t jb_plan_enumerate(int32_t m, int32_t p, int32_t rank, int32_t rloc, int64_t ldw, int32_t t,
                          int32_t piv_rank, int32_t piv_slot, int32_t do_swap, int32_t phase,
                          int32_t pct0, int32_t pct1, int32_t nb1, int32_t nearly,
                          int32_t next_piv_rank, int32_t next_piv_slot, int32_t* out, int64_t cap) {
  StepParams P;
  memset(&P, 0, sizeof(P));
  P.m = m; P.p = p; P.rank = rank; P.t = t; P.rloc = rloc; P.ldw = ldw;
  P.nsub = (m + JB_BM - 1) / JB_BM;
  P.CT = (int)((ldw + JB_BN - 1) / JB_BN);
  P.t_slot = (t % p == rank) ? t / p : -1;
  P.phase = phase; P.pct0 = pct0; P.pct1 = pct1; P.nb1 = nb1; P.nearly = nearly;
  P.has_next = (phase == JB_PHASE_B2 && next_piv_rank >= 0) ? 1 : 0;
  P.sc_t = (phase == JB_PHASE_B2) ? t + 1 : t;
  JbStepInfo si;
  memset(&si, 0, sizeof(si));
  si.piv_rank = piv_rank; si.piv_slot = piv_slot; si.do_swap = do_swap; si.piv_g = piv_slot * p + piv_rank;
  JbStepInfo sn;
  memset(&sn, 0, sizeof(sn));
  sn.piv_rank = next_piv_rank; sn.piv_slot = next_piv_slot; sn.piv_g = next_piv_slot * p + next_piv_rank;
  const jbplan::Plan pl = jbplan::make_plan(P, si, sn);
  if (pl.total > cap) return -1;
  for (int idx = 0; idx < pl.total; ++idx) {
    const jbplan::Item it = jbplan::decode(P, si, sn, pl, idx);
    out[5 * idx] = it.type; out[5 * idx + 1] = it.ct; out[5 * idx + 2] = it.lrow0; out[5 * idx + 3] = it.rows_valid;
    out[5 * idx + 4] = it.prow;
  }
  return pl.total;
}

}  // extern "C"
