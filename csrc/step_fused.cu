// Fused elimination-step kernel (K2 `pivot_row_scale_mcast` + K3 `trailing_update_f64` + K4 row swap).
//
// One persistent launch per step and rank.  CTAs walk a common work list:
//   SCALE items (pivot owner only): R_j = P^-1 * W(piv, j) on FP64 tensor cores, result written
//       straight into EVERY rank's pivot-row buffer (NVLS multimem.st, or unicast peer stores when
//       no multicast object is bound), then a per-column-tile flag is bumped on every rank.
//   ROWT items (owner of block row t): push the old row t into the pivot owner's swap buffer over
//       NVLink (the block-row swap of main.cpp:1100-1131 done with peer stores instead of
//       MPI_Send/Recv), then replace row t by the scaled pivot row.
//   UPD  items (every rank): C(i,j) += A(i,t) * (-R_j) for all local rows, DMMA m8n8k4, operands fed by
//       TMA (cp.async.bulk.tensor) through a 4-stage mbarrier ring.  A consumer tile starts as soon
//       as the flag of ITS column tile has landed, so broadcast and math overlap tile by tile.
//
// Reference semantics (main.cpp:1132-1194): pivot row <- P^-1 * row; every other block row i:
// row_i -= A(i,t) * pivot_row.  We run the in-place form of Gauss-Jordan: the B = I half of
// [A|B] is never stored; column slot t receives P^-1 (pivot row) or -A(i,t)*P^-1 (other rows),
// which is exactly what the reference leaves in column pi(t) of B.  Work drops from the
// reference's 3n^3 to 2n^3 flops with identical arithmetic on the non-zero terms.
#include "common.cuh"
#include "engine_types.h"
#include "kernels.h"
#include "step_plan.h"

using namespace jb;

namespace {

constexpr int MMA_WARPS = 4;
constexpr int STEP_THREADS = (MMA_WARPS + 1) * 32;  // + 1 TMA producer warp
constexpr int L_BYTES = JB_LS * JB_BM * 8;          // 10240
constexpr int R_BYTES = JB_RS * JB_BK * 8;          // 16896
constexpr int STAGE_BYTES = L_BYTES + R_BYTES;      // 27136
constexpr int RING = 4;                             // depth of the claimed-item ring

using namespace jbplan;

__device__ __forceinline__ double2 ld_f64x2(const double* p) {
  return *reinterpret_cast<const double2*>(p);
}
__device__ __forceinline__ void st_f64x2(double* p, double x, double y) {
  *reinterpret_cast<double2*>(p) = make_double2(x, y);
}

// ---- K4: swap + install the scaled pivot row as block row t (one 64-row chunk per item).
//      Thread -> (column pair v, row parity); 4 independent 16-byte loads in flight per thread.
//      Not inlined: its address arithmetic must not compete for registers with the DMMA loop.
__device__ __noinline__ void rowt_item(double* wrow, double* swap_dst, uint32_t* swap_flag,
                                       const uint32_t* dest_done, uint32_t dest_want,
                                       const double* src, const uint32_t* row_flag, uint32_t row_expect,
                                       uint32_t tag, int rows_valid, int ncol, int64_t ldw,
                                       int* abort_flag, unsigned long long spin_to) {
  const int tid = threadIdx.x, lane = tid & 31;
  const int vec_per_row = ncol / 2;
  const int v = tid & 63, rh = tid >> 6;
  const bool vok = v < vec_per_row;
  wrow += 2 * v;
  src += 2 * v;
  if (swap_dst) {
    // the old block row t goes to the pivot owner's swap buffer over NVLink (main.cpp:1100-1131);
    // the buffer (parity t & 1) was last used by step t-2: the owner must have retired that step
    if (dest_done) {
      if (lane == 0) wait_flag_ge(dest_done, dest_want, abort_flag, spin_to);
      __syncwarp();
    }
    double* dst = swap_dst + 2 * v;
    for (int r0 = rh; r0 < rows_valid; r0 += 8) {
      double2 x[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        x[u] = make_double2(0.0, 0.0);
        if (vok && r0 + 2 * u < rows_valid) x[u] = ld_f64x2(wrow + (int64_t)(r0 + 2 * u) * ldw);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (vok && r0 + 2 * u < rows_valid) st_f64x2(dst + (int64_t)(r0 + 2 * u) * ldw, x[u].x, x[u].y);
    }
    __threadfence_system();
    named_bar_sync(1, MMA_WARPS * 32);
    if (tid == 0) st_release_sys_u32(swap_flag, tag);
  }
  if (lane == 0) wait_flag_ge(row_flag, row_expect, abort_flag, spin_to);
  __syncwarp();
  for (int r0 = rh; r0 < rows_valid; r0 += 8) {
    double2 x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      x[u] = make_double2(0.0, 0.0);
      if (vok && r0 + 2 * u < rows_valid) x[u] = ld_f64x2(src + (int64_t)(r0 + 2 * u) * ldw);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (vok && r0 + 2 * u < rows_valid)
        st_f64x2(wrow + (int64_t)(r0 + 2 * u) * ldw, -x[u].x, -x[u].y);  // rowbuf holds -R
  }
}

// B1's early column tiles: every item of the tile (UPD of each row tile, ROWT of block row t) calls
// this after its stores; the last one flags the tile as carrying step t.  All 4 MMA warps.
__device__ __forceinline__ void early_arrive(const StepParams& P, const Plan& pl, int epos) {
  __threadfence();
  named_bar_sync(1, MMA_WARPS * 32);
  if (threadIdx.x == 0) {
    const uint32_t want = (uint32_t)(pl.rtu + (P.t_slot >= 0 ? P.nsub : 0));
    const uint32_t old = atomicAdd(P.early + epos, 1u);
    if (old == want - 1) {
      P.early[epos] = 0;
      st_release_sys_u32(P.early + 8 + epos, P.tag);
    }
  }
}

// One thread per retiring CTA of phase B1; the last one publishes "B1(t) done".
__device__ __forceinline__ void b1_retire(const StepParams& P) {
  const uint32_t old = atomicAdd(P.b1sync, 1u);
  if (old == gridDim.x - 1) {
    P.b1sync[0] = 0;
    st_release_sys_u32(P.b1sync + 1, P.tag);
  }
}
// ... and of phase B2: the last one waits for B1(t) (an earlier launch that never waits for B2) and
// tells every rank that this rank has retired the bulk of step t.
__device__ __forceinline__ void b2_retire(const StepParams& P) {
  const uint32_t old = atomicAdd(P.b1sync + 2, 1u);
  if (old == gridDim.x - 1) {
    P.b1sync[2] = 0;
    wait_flag_ge(P.b1sync + 1, P.tag, const_cast<int*>(&P.ctl->timeout_flag), P.ctl->spin_timeout_ns);
    __threadfence_system();
    for (int q = 0; q < P.p; ++q) st_release_sys_u32(P.stepdone_peers[q] + P.rank, P.tag);
  }
}

// kMode selects which item kinds a launch can contain (bit 0: SCALE, bit 1: ROWT; UPD always), so
// that code a phase never runs is compiled out and cannot disturb the register allocation of the
// DMMA loop:  0 = B (tcgen05 mode), 1 = S / B2, 2 = A / B1, 3 = ALL / A2.
template <int kMode>
__global__ void __launch_bounds__(STEP_THREADS, 2)
k_step_fused(const StepParams P, const __grid_constant__ StepMaps M) {
  constexpr bool kScale = (kMode & 1) != 0;
  constexpr bool kRowt = (kMode & 2) != 0;
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full_bar[JB_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[JB_STAGES];
  // work ring: the producer thread claims items from the global counter and hands them to the MMA
  // warps (dynamic scheduling: a CTA that starts late, e.g. after a K1 CTA left its SM, simply
  // takes whatever work is left)
  __shared__ __align__(8) uint64_t ring_full[RING];
  __shared__ __align__(8) uint64_t ring_empty[RING];
  __shared__ int ring_item[RING];

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;

  if (P.ctl->singular | P.ctl->timeout_flag) return;
  int* const abort_flag = const_cast<int*>(&P.ctl->timeout_flag);
  const unsigned long long spin_to = P.ctl->spin_timeout_ns;
  const JbStepInfo si = P.stepinfo[P.t];
  // phase B2 works for two steps at once: UPD items belong to step t, SCALE items to step t + 1
  const JbStepInfo sn = (kScale && P.phase == JB_PHASE_B2 && P.has_next) ? P.stepinfo[P.t + 1] : si;
  const Plan pl = make_plan(P, si, sn);
  if (pl.total == 0) {
    if (kRowt && !kScale && P.phase == JB_PHASE_B1 && P.b1sync && tid == 0) b1_retire(P);
    if (kScale && !kRowt && P.phase == JB_PHASE_B2 && P.b1sync && tid == 0) b2_retire(P);
    return;
  }
  const int sc_slot = (P.sc_t == P.t) ? si.piv_slot : sn.piv_slot;  // pivot row the SCALE items read

#ifdef JB_DBG_GENERIC
  unsigned char* stage_base =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem) + 127) & ~uintptr_t(127));
#else
  unsigned char* const stage_base = smem;  // TMA destinations need 128-byte alignment
#endif

  if (tid == 0) {
    for (int s = 0; s < JB_STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], MMA_WARPS);
    }
    for (int s = 0; s < RING; ++s) {
      mbar_init(&ring_full[s], 1);
      mbar_init(&ring_empty[s], MMA_WARPS);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int KC = P.Kp / JB_BK;
  const int col_t0 = P.t * P.m, col_t1 = col_t0 + P.m;  // column slot of step t
  const int sc_col_t0 = P.sc_t * P.m, sc_col_t1 = sc_col_t0 + P.m;  // ... of the SCALE items' step
  double* const rowbuf_local = P.rowbuf_peers[P.rank];
  const uint32_t* const rowflag_local = P.rowflag_peers[P.rank];
  double* const swapbuf_local = P.swapbuf_peers[P.rank];
  const uint32_t* const swapflag_local = P.swapflag_peers[P.rank];
  // column tiles holding the swapped-in row's leading block A(t,t)
  const int ctl0 = col_t0 / JB_BN, ctl1 = (col_t1 - 1) / JB_BN;

  if (warp == MMA_WARPS) {
    // =============================================================== TMA producer warp
    if (lane == 0) {
      tma_prefetch_desc(&M.tmW);
      tma_prefetch_desc(&M.tmRow);
      tma_prefetch_desc(&M.tmL);
      tma_prefetch_desc(&M.tmInv);
      tma_prefetch_desc(&M.tmSwap);
      uint32_t it_cnt = 0;
      bool peers_done = false;
      for (uint32_t rp = 0;; ++rp) {
        const int idx = (int)atomicAdd(P.counter, 1u);
        const int slot = rp % RING;
        mbar_wait(&ring_empty[slot], ((rp / RING) & 1) ^ 1);
        ring_item[slot] = idx;
        mbar_arrive(&ring_full[slot]);  // release: publishes ring_item[slot]
        if (idx >= pl.total) break;
        const Item it = decode(P, si, sn, pl, idx);
        if (kRowt && it.type == IT_ROWT) continue;
        if (P.only_swapped && it.type != IT_UPD_SWAPPED) continue;  // tcgen05 mode: the rest is done there
        const int col0 = it.ct * JB_BN;
        if (!kScale || it.type != IT_SCALE) {
          if (kRowt && P.phase == JB_PHASE_A && (it.ct < P.prev_pct0 || it.ct > P.prev_pct1)) {
            // phase A runs beside the bulk of the previous step: this tile is one of B1(t-1)'s
            // early tiles, flagged when every row carries update t-1.  NOTE: this wait orders the
            // TMA operand loads only; the MMA warps read the C tile of the item right after the
            // hand-off.  That is fine because the engine launches phase A behind a FINISHED B1(t-1)
            // (the flag is then already set); with JB_B2_OVERLAP the MMA warps would have to wait too.
            int j = it.ct - P.prev_pct1 - 1;
            if (j < 0) j += P.CT;
            if (j < 8) wait_flag_ge(P.early + 8 + j, P.prev_tag, abort_flag, spin_to);
          }
          wait_flag_ge(rowflag_local + it.ct, P.row_expect, abort_flag, spin_to);  // pivot-row tile landed
          if (it.type == IT_UPD_SWAPPED)
            for (int c = ctl0; c <= ctl1; ++c) wait_flag_ge(swapflag_local + c * P.nsub + it.sub, P.tag, abort_flag, spin_to);
          fence_proxy_async();  // remote generic-proxy writes -> my async-proxy (TMA) reads
        } else if (P.phase == JB_PHASE_B2) {
          // SCALE(t+1) reads the next pivot row as left by UPD(t).  The tiles of THIS phase are the
          // UPDP items at the head of the work list (lower claim index, hence running): wait for
          // their flags.  Priority tiles were finished by phase A (an earlier launch on the way to
          // K1(t+1)).  The rest belongs to phase B1, which (JB_B2_OVERLAP only) may still be running but never
          // waits for this launch: its first `nearly` tiles are flagged one by one, for the others
          // wait until its last CTA has retired.
          if (!peers_done && P.t > 0) {
            // rowbuf / rowflag of step t+1 (parity (t+1) & 1) were last read by step t-1: every rank
            // must have retired it before the first tile is overwritten
            for (int q = 0; q < P.p; ++q)
              wait_flag_ge(P.stepdone_peers[P.rank] + q, P.tag - 1, abort_flag, spin_to);
            peers_done = true;
          }
          const bool pri = pl.npri > 0 && it.ct >= P.pct0 && it.ct <= P.pct1;
          if (!pri) {
            const int pos = np_pos(P, pl, it.ct);
            if (pos >= pl.c0) {
              for (int sb = 0; sb < P.nsub; ++sb)
                wait_flag_ge(P.prowflag + it.ct * P.nsub + sb, P.tag, abort_flag, spin_to);
            } else if (pos < P.nearly) {
              wait_flag_ge(P.early + 8 + pos, P.tag, abort_flag, spin_to);  // B1 flags these early
            } else if (P.b1sync) {
              wait_flag_ge(P.b1sync + 1, P.tag, abort_flag, spin_to);
            }
          }
          fence_proxy_async();  // other CTAs' generic-proxy stores -> my TMA reads of W
        }
        for (int c = 0; c < KC; ++c, ++it_cnt) {
          const int s = it_cnt % JB_STAGES;
          const uint32_t ph = (it_cnt / JB_STAGES) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          unsigned char* sb = stage_base + (size_t)s * STAGE_BYTES;
          mbar_expect_tx(&full_bar[s], STAGE_BYTES);
          const int k0 = c * JB_BK;
          if (kScale && it.type == IT_SCALE) {
            tma_load_2d(sb, &M.tmInv, k0, sc_slot * P.Kp + it.sub * JB_BM, &full_bar[s]);
            tma_load_2d(sb + L_BYTES, &M.tmW, col0, sc_slot * P.m + k0, &full_bar[s]);
          } else if (it.type == IT_UPD) {
            tma_load_2d(sb, &M.tmL, k0, it.lrow0, &full_bar[s]);
            tma_load_2d(sb + L_BYTES, &M.tmRow, col0, k0, &full_bar[s]);
          } else {
            tma_load_2d(sb, &M.tmSwap, col_t0 + k0, it.sub * JB_BM, &full_bar[s]);
            tma_load_2d(sb + L_BYTES, &M.tmRow, col0, k0, &full_bar[s]);
          }
        }
      }
    }
    return;
  }

  // ================================================================= MMA / copy warps
  const int g = lane >> 2, tig = lane & 3;
  const int wn = warp;  // 4 warps side by side along N, each 64 x 32
  uint32_t it_cnt = 0;

  for (uint32_t rc = 0;; ++rc) {
    const int rslot = rc % RING;
    mbar_wait(&ring_full[rslot], (rc / RING) & 1);
    const int idx = ring_item[rslot];
    __syncwarp();
    if (lane == 0) mbar_arrive(&ring_empty[rslot]);
    if (idx >= pl.total) break;
    const Item it = decode(P, si, sn, pl, idx);
    const int col0 = it.ct * JB_BN;
    if (P.only_swapped && it.type != IT_UPD_SWAPPED) continue;

    if (kRowt && it.type == IT_ROWT) {
      if (P.phase == JB_PHASE_A && (it.ct < P.prev_pct0 || it.ct > P.prev_pct1)) {
        int j = it.ct - P.prev_pct1 - 1;
        if (j < 0) j += P.CT;
        if (j < 8 && lane == 0) wait_flag_ge(P.early + 8 + j, P.prev_tag, abort_flag, spin_to);
        __syncwarp();
      }
      const int ncol = (int)min((int64_t)JB_BN, P.ldw - col0);  // multiple of 16
      const int64_t roff = (int64_t)(it.sub * JB_BM) * P.ldw + col0;
      rowt_item(P.W + (int64_t)P.t_slot * P.m * P.ldw + roff,
                si.do_swap ? P.swapbuf_peers[si.piv_rank] + roff : nullptr,
                si.do_swap ? P.swapflag_peers[si.piv_rank] + it.ct * P.nsub + it.sub : nullptr,
                (P.b1sync && P.t >= 2) ? P.stepdone_peers[P.rank] + si.piv_rank : nullptr, P.tag - 2,
                rowbuf_local + roff, rowflag_local + it.ct, P.row_expect, P.tag, it.rows_valid, ncol,
                P.ldw, abort_flag, spin_to);
      // B1's early column tiles: block row t is part of "this tile carries step t" - phase A of the
      // next step updates row t as an ordinary row and K1(t+2) copies it into the L panel - so the
      // install counts towards the tile's flag like the UPD items do
      if (!kScale && P.phase == JB_PHASE_B1) {
        const int epos = np_pos(P, pl, it.ct);
        if (epos < P.nearly) early_arrive(P, pl, epos);
      }
      continue;
    }

    // ---- accumulator init
    double acc[8][4][2];
    const bool upd = !kScale || (it.type != IT_SCALE);
    if (kScale && it.type == IT_SCALE) {
#pragma unroll
      for (int mt = 0; mt < 8; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;
    } else {
      const double* cbase;
      if (it.type == IT_UPD) {
        cbase = P.W + (int64_t)it.lrow0 * P.ldw;
      } else {
        if (lane == 0) {
          wait_flag_ge(swapflag_local + it.ct * P.nsub + it.sub, P.tag, abort_flag, spin_to);
        }
        __syncwarp();
        cbase = swapbuf_local + (int64_t)(it.sub * JB_BM) * P.ldw;
      }
#pragma unroll
      for (int mt = 0; mt < 8; ++mt) {
        const int r = mt * 8 + g;
        const bool rok = r < it.rows_valid;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const int c = col0 + wn * 32 + nt * 8 + 2 * tig;
          double2 v = make_double2(0.0, 0.0);
          if (rok && c < P.ldw) v = ld_f64x2(cbase + (int64_t)r * P.ldw + c);
          // column slot t starts from the identity column of B, i.e. zero for non-pivot rows
          if (c >= col_t0 && c < col_t1) v.x = 0.0;
          if (c + 1 >= col_t0 && c + 1 < col_t1) v.y = 0.0;
          acc[mt][nt][0] = v.x;
          acc[mt][nt][1] = v.y;
        }
      }
    }

    // ---- main loop over K = m in chunks of JB_BK, DMMA m8n8k4
    for (int c = 0; c < KC; ++c, ++it_cnt) {
      const int s = it_cnt % JB_STAGES;
      const uint32_t ph = (it_cnt / JB_STAGES) & 1;
      mbar_wait(&full_bar[s], ph);
      const double* Ls = reinterpret_cast<const double*>(stage_base + (size_t)s * STAGE_BYTES);
      const double* Rs = Ls + JB_LS * JB_BM;
#pragma unroll
      for (int kk = 0; kk < JB_BK / 4; ++kk) {
        double a[8], b[4];
#pragma unroll
        for (int mt = 0; mt < 8; ++mt) {
          a[mt] = Ls[(mt * 8 + g) * JB_LS + kk * 4 + tig];
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) b[nt] = Rs[(kk * 4 + tig) * JB_RS + wn * 32 + nt * 8 + g];
#pragma unroll
        for (int mt = 0; mt < 8; ++mt)
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], b[nt]);
      }
      // Stage release is a WAR hazard ACROSS PROXIES: our LDS reads (generic proxy) must be ordered
      // before the producer's next TMA write (async proxy) into the same stage.  The mbarrier
      // arrive alone only orders the generic proxy (ptxas even hoists it above the consuming
      // DMMAs), which showed up as sporadically corrupted tiles once W exceeded L2.
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    // ---- epilogue
    if (upd) {
      double* obase = P.W + (int64_t)it.lrow0 * P.ldw;
#pragma unroll
      for (int mt = 0; mt < 8; ++mt) {
        const int r = mt * 8 + g;
        if (r < it.rows_valid) {
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {
            const int c = col0 + wn * 32 + nt * 8 + 2 * tig;
            if (c < P.ldw) st_f64x2(obase + (int64_t)r * P.ldw + c, acc[mt][nt][0], acc[mt][nt][1]);
          }
        }
      }
      if (kScale && it.prow) {  // this tile of the next pivot row now carries update t: release SCALE(t+1)
        __threadfence();
        named_bar_sync(1, MMA_WARPS * 32);
        if (tid == 0) st_release_sys_u32(P.prowflag + it.ct * P.nsub + it.sub, P.tag);
      }
      if (kRowt && !kScale && it.epos >= 0) early_arrive(P, pl, it.epos);  // B1's early column tiles
    } else if (kScale) {
      // SCALE: column slot t of the pivot row is P^-1 itself; broadcast the tile to all ranks
      const double* pinv = P.invbuf + (int64_t)sc_slot * P.Kp * P.Kp;
#pragma unroll
      for (int mt = 0; mt < 8; ++mt) {
        const int r = it.sub * JB_BM + mt * 8 + g;  // row inside the pivot block
        if (mt * 8 + g < it.rows_valid) {
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {
            const int c = col0 + wn * 32 + nt * 8 + 2 * tig;
            if (c < P.ldw) {
              // the buffer holds -R so the update is a plain accumulate C += L * (-R)
              double x = -acc[mt][nt][0], y = -acc[mt][nt][1];
              if (c >= sc_col_t0 && c < sc_col_t1) x = -pinv[(int64_t)r * P.Kp + (c - sc_col_t0)];
              if (c + 1 >= sc_col_t0 && c + 1 < sc_col_t1) y = -pinv[(int64_t)r * P.Kp + (c + 1 - sc_col_t0)];
              const int64_t off = (int64_t)r * P.ldw + c;
              if (P.sc_rowbuf_mc) {
                multimem_st_v2f64(P.sc_rowbuf_mc + off, x, y);
              } else {
#pragma unroll 1
                for (int q = 0; q < P.p; ++q) st_f64x2(P.sc_rowbuf_peers[q] + off, x, y);
              }
            }
          }
        }
      }
      __threadfence_system();
      named_bar_sync(1, MMA_WARPS * 32);
      if (tid == 0) {
        if (P.sc_rowflag_mc) {
          multimem_red_release_add_u32(P.sc_rowflag_mc + it.ct, 1u);
        } else {
#pragma unroll 1
          for (int q = 0; q < P.p; ++q) red_release_sys_add_u32(P.sc_rowflag_peers[q] + it.ct, 1u);
        }
      }
    }
  }
  if (kRowt && !kScale && P.phase == JB_PHASE_B1 && P.b1sync) {
    // B1 only: the last CTA to retire tells B2 (possibly already running beside us) that every
    // tile of this phase carries update t
    __threadfence();
    named_bar_sync(1, MMA_WARPS * 32);
    if (tid == 0) b1_retire(P);
  }
  if (kScale && !kRowt && P.phase == JB_PHASE_B2 && P.b1sync) {
    __threadfence();
    named_bar_sync(1, MMA_WARPS * 32);
    if (tid == 0) b2_retire(P);
  }
}

}  // namespace

size_t jb_step_smem_bytes() { return (size_t)JB_STAGES * STAGE_BYTES; }

template <int kMode>
static cudaError_t launch_mode(const StepParams& P, const StepMaps& M, int grid, cudaStream_t st) {
  static bool configured[JB_MAX_RANKS] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev % JB_MAX_RANKS]) {
    cudaError_t e = cudaFuncSetAttribute(k_step_fused<kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)jb_step_smem_bytes());
    if (e != cudaSuccess) return e;
    configured[dev % JB_MAX_RANKS] = true;
  }
  k_step_fused<kMode><<<grid, STEP_THREADS, jb_step_smem_bytes(), st>>>(P, M);
  return cudaGetLastError();
}

cudaError_t jb_launch_step(const StepParams& P, const StepMaps& M, int grid, cudaStream_t st) {
  switch (P.phase) {
    case JB_PHASE_B: return launch_mode<0>(P, M, grid, st);
    case JB_PHASE_S:
    case JB_PHASE_B2: return launch_mode<1>(P, M, grid, st);
    case JB_PHASE_A:
    case JB_PHASE_B1: return launch_mode<2>(P, M, grid, st);
    default: return launch_mode<3>(P, M, grid, st);
  }
}
