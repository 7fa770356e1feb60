// Kernel parameter blocks and host-side launchers (internal to libjordan_b200).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "engine_types.h"

// ------------------------------------------------------------------ K1 pivot search + min-loc
struct K1Params {
  const double* W;     // local rows, ld = ldw
  double* Lbuf;        // [rloc*m][Kp] copy of block-column t
  double* invbuf;      // [rloc][Kp][Kp] inverse of each candidate
  double* normbuf;     // [rloc]
  int* okbuf;          // [rloc]
  JbCtl* ctl;
  JbStepInfo* stepinfo;
  JbMail* mail_peers[JB_MAX_RANKS];  // every rank's mailbox ([2][JB_MAX_RANKS]); [rank] is local
  int64_t ldw;
  int m, Kp, p, rank, t;
  int rloc;     // local block rows
  int s0;       // first candidate slot (global index >= t)
  int ncand;    // number of candidates on this rank (may be 0)
  int ngroups;  // max(ncand, 1) CTA groups doing inversion / min-loc
  uint32_t tag; // global step counter + 1
};
int jb_k1_cluster_size(int m);
cudaError_t jb_launch_k1(const K1Params& P, int panel_ctas, cudaStream_t st);

// ------------------------------------------------------------------ fused step kernel (K2+K3+K4)
struct StepParams {
  double* W;
  const double* invbuf;
  const JbStepInfo* stepinfo;
  const JbCtl* ctl;
  double* rowbuf_peers[JB_MAX_RANKS];     // rowbuf[par(t)] on every rank
  uint32_t* rowflag_peers[JB_MAX_RANKS];  // rowflag[par(t)] on every rank
  double* swapbuf_peers[JB_MAX_RANKS];    // swapbuf[par(t)] on every rank
  uint32_t* swapflag_peers[JB_MAX_RANKS];
  double* rowbuf_mc;      // NVLS multicast alias of rowbuf[par(t)] (nullptr -> unicast peer stores)
  uint32_t* rowflag_mc;   // NVLS multicast alias of rowflag[par(t)]
  // SCALE context: the step whose pivot row the SCALE items of this launch produce.  sc_t == t in
  // the serial schedule and the step-0 prologue; sc_t == t + 1 in phase B2 (pivot-row look-ahead:
  // the owner of the NEXT pivot row scales and broadcasts it while step t is still being applied).
  double* sc_rowbuf_peers[JB_MAX_RANKS];
  uint32_t* sc_rowflag_peers[JB_MAX_RANKS];
  double* sc_rowbuf_mc;
  uint32_t* sc_rowflag_mc;
  int sc_t;
  uint32_t* prowflag;     // [CT][nsub] local: tag of the last step whose update reached this tile of
                          // the NEXT pivot row (written by the UPDP items of phase B2)
  uint32_t* b1sync;       // local {CTA exit counter, done tag}: the last CTA of phase B1(t) to retire
                          // publishes tag(t); B2(t), which may run beside B1(t), waits for it before
                          // SCALE(t+1) touches a column tile B1 updates.  nullptr = not used
  int64_t ldw;
  int m, Kp, p, rank, t;
  int rloc, nsub, CT;
  int t_slot;             // local slot of block row t on its owner, else -1
  uint32_t row_expect;    // rowflag value meaning "tile landed"
  uint32_t tag;           // swapflag value for this step
  int phase;              // JB_PHASE_*
  int pct0, pct1;         // priority column tiles (cover block column t+1); empty if pct1 < pct0
  // Cross-rank "step finished" tags (symmetric region): stepdone_peers[r][q] on rank r holds the tag
  // of the last step whose bulk update rank q has completely retired (B1 and B2).  With phase A and
  // K1 running ahead of the bulk, the min-loc exchange no longer implies that every rank is done
  // with step t-2, so whoever overwrites a parity-double-buffered peer buffer of step t (SCALE into
  // rowbuf/rowflag, ROWT push into the pivot owner's swapbuf) first waits for tag(t-2) there.
  uint32_t* stepdone_peers[JB_MAX_RANKS];
  int prev_pct0, prev_pct1;  // phase A: priority range of the previous step; tiles outside it are
  uint32_t prev_tag;         // B1(t-1)'s early tiles, complete once early[8 + j] >= prev_tag
  int nb1;                // look-ahead: the first nb1 non-priority column tiles belong to phase B1
  int nearly;             // ... and B1 flags each of its first `nearly` column tiles when complete (<= 8)
  uint32_t* early;        // local [16]: [0..7] item counters, [8..15] done tags of those tiles
  int has_next;           // phase B2: stepinfo[t + 1] is valid (t + 1 < Nr)
  uint32_t* counter;      // work-item counter of THIS launch (zeroed by the host)
  int only_swapped;       // phase B in tcgen05 mode: only the swapped-in block row stays on the fp64 path
};
// Work-list phases (csrc/step_plan.h):
//   ALL  serial schedule: SCALE(t) + ROWT(t) + UPD(t) over every column tile
//   S    SCALE(t) over every column tile (prologue of the look-ahead schedule, t = 0)
//   A    ROWT(t) + UPD(t) of the priority column tiles (what K1(t+1) needs)
//   B1   ROWT(t) of the remaining tiles + UPD(t) of the first nb1 of them (K1(t+1) runs beside it)
//   B2   after K1(t+1): UPD(t) of the next pivot row first, SCALE(t+1) of every tile, rest of UPD(t)
//   A2/B tcgen05 mode: SCALE + ROWT of every tile / fp64 update of the swapped-in row only
enum { JB_PHASE_ALL = 0, JB_PHASE_A = 1, JB_PHASE_B = 2, JB_PHASE_A2 = 3, JB_PHASE_S = 4,
       JB_PHASE_B1 = 5, JB_PHASE_B2 = 6 };
struct StepMaps {
  CUtensorMap tmW;     // W              box {JB_RS, JB_BK}
  CUtensorMap tmRow;   // rowbuf[par]    box {JB_RS, JB_BK}
  CUtensorMap tmL;     // Lbuf           box {JB_LS, JB_BM}
  CUtensorMap tmInv;   // invbuf         box {JB_LS, JB_BM}
  CUtensorMap tmSwap;  // swapbuf        box {JB_LS, JB_BM}
};
size_t jb_step_smem_bytes();
cudaError_t jb_launch_step(const StepParams& P, const StepMaps& M, int grid, cudaStream_t st);

// ------------------------------------------------------------------ K5 generator fill + norm
struct FillParams {
  double* W;
  int64_t ldw, n;
  int m, p, rank, rloc, formula;
  double* rowsum_max;  // device double, atomic max of abs row sums (local ||A||inf)
};
cudaError_t jb_launch_fill(const FillParams& P, cudaStream_t st);
cudaError_t jb_launch_rowsum_max(const double* W, int64_t rows, int64_t cols, int64_t ld,
                                 double* out_max, cudaStream_t st);

// blocks with pitch Kp -> packed m x m blocks (stand-alone block inversion entry points)
cudaError_t jb_launch_compact_blocks(const double* src, int Kp, double* dst, int m, int64_t nblocks,
                                     cudaStream_t st);

// ------------------------------------------------------------------ final column un-permutation
cudaError_t jb_launch_unpermute(double* W, int64_t rows, int64_t ldw, int m, int Nr,
                                const JbStepInfo* stepinfo, cudaStream_t st);

// ------------------------------------------------------------------ K6/K7 distributed GEMM
// C_loc (rows_loc x n) = A_loc (rows_loc x n; formula-generated or stored) * X (n x n, block-cyclic
// rows spread over the ranks).  A column panel of X is first pulled from the owning peers' HBM
// into a local staging panel (k_gather_panel), then multiplied on the FP64 tensor cores.
enum { JB_GEMM_RESIDUAL = 0, JB_GEMM_STORE_R = 1, JB_GEMM_XPLUS = 2 };
struct GatherParams {
  const double* X_peers[JB_MAX_RANKS];  // every rank's local slab (block-cyclic rows), ld = ldx
  int64_t ldx;
  double* Xp;                           // local panel [n][ldxp], rows in GLOBAL order
  int64_t ldxp;
  int64_t n, col0;
  int m, p, pw;
};
struct DistGemmParams {
  const double* A;      // stored left operand, local rows (nullptr -> generator formula)
  int64_t lda;
  int formula;
  const double* Xp;     // gathered right-operand panel [n][ldxp]
  int64_t ldxp;
  int64_t col0;         // first global column covered by the panel
  int pw;               // panel width
  double* C;            // output slab (STORE_R / XPLUS)
  int64_t ldc;
  const double* Cin;    // XPLUS: C = Cin + A*Xp
  int64_t ldcin;
  double* rowsum;       // RESIDUAL: [rows_loc] abs row sums, accumulated with atomics
  int64_t n;
  int m, p, rank, rloc, mode;
};
cudaError_t jb_launch_gather_panel(const GatherParams& P, cudaStream_t st);
cudaError_t jb_launch_dist_gemm(const DistGemmParams& P, cudaStream_t st);
cudaError_t jb_launch_max_reduce(const double* v, int64_t count, double* out, cudaStream_t st);

// ------------------------------------------------------------------ K3b tcgen05 int8-split update
struct Tc05Launch {
  double* C;
  int64_t ldc, rows, cols;
  const int8_t* Lq;       // [S][KB][rows_pad][128]
  const int8_t* Rq;       // [S][KB][cols_pad][128]
  int64_t rows_pad, cols_pad;
  const double* rscale;   // [rows]
  const double* cscale;   // [cols_pad]
  int K, S, grid;
  int64_t skip0, skiplen; // local rows handled elsewhere (block row t)
  const JbCtl* ctl;
  const JbStepInfo* stepinfo;
  int t, m, rank;
  int col_t0, col_t1;     // column slot t (accumulate from zero), -1/-1 = none
  int zero_slot;          // v2: clear column slot t before this launch (first k-block only)
};
size_t jb_tc05_smem_bytes(int S);
size_t jb_tc05_lq_bytes(int64_t rows_pad, int K, int S);
size_t jb_tc05_rq_bytes(int64_t cols_pad, int K, int S);
cudaError_t jb_tc05_slice_rows(const double* L, int64_t ldl, int64_t rows, int K, int S, int8_t* Lq,
                               int64_t rows_pad, double* scale, cudaStream_t st);
cudaError_t jb_tc05_slice_cols(const double* Rn, int64_t ldr, int64_t cols, int K, int S, int8_t* Rq,
                               int64_t cols_pad, double* scale, const uint32_t* flags, uint32_t expect,
                               const JbCtl* ctl, cudaStream_t st);
int jb_tc05_launch(const Tc05Launch& a, cudaStream_t st);

#include "kernels_formula.h"
