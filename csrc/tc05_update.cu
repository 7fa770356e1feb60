// K3b `trailing_update_tc05` — the trailing update C += L * (-R) on the 5th-generation tensor cores
// (tcgen05.mma, accumulators in TMEM) with an error-free integer split of the fp64 operands
// (Ozaki-style), optional path selected by JbConfig.precision = 1.
//
// Why integers and not a tf32 / bf16 split with fp32 accumulation: kappa_inf(|i-j|) ~ 1e9 at
// n = 32768, so an fp32-accumulated elimination (eps ~ 6e-8) does not even give a convergent Newton
// start.  tcgen05.mma kind::i8 accumulates exactly in int32, so with 7-bit signed digits
//     L(i,k)  = 2^eL(i) * sum_s qL_s(i,k) 2^(-6-7s),   |qL_s| <= 64   (row scaling)
//     -R(k,j) = 2^eR(j) * sum_u qR_u(k,j) 2^(-6-7u),   |qR_u| <= 64   (column scaling)
// every slice product sum_k qL_s qR_u is exact, the pairs with s+u = d share one TMEM accumulator
// (|D_d| < 2^23), and  L*(-R) = 2^(eL+eR-12) * sum_d D_d 2^(-7d)  is recombined in fp64 by the
// epilogue.  S = 8 slices keep 55 bits below each row / column maximum, i.e. fp64-level accuracy.
//
// Kernel layout (1 CTA / SM, 192 threads): warp 0 = TMA producer (SWIZZLE_128B K-major int8 tiles),
// warp 1 = TMEM allocator + single-thread MMA issuer (36 slice pairs x 4 K-steps per 128x64 tile),
// warps 2-5 = epilogue (tcgen05.ld -> Horner recombination in fp64 -> C read-modify-write).
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <mutex>
#include <vector>
#include "common.cuh"
#include "engine_types.h"
#include "kernels.h"
#include "../include/jordan_b200.h"

using namespace jb;

namespace {

constexpr int TC_M = 128;      // rows per tile (UMMA M)
constexpr int TC_N = 64;       // columns per tile (UMMA N); S accumulators x 64 columns = 512 TMEM columns
constexpr int TC_K = 128;      // K per k-block = one 128-byte swizzle atom of int8
constexpr int TC_SMAX = 8;     // slices
constexpr int TC_THREADS = 192;
constexpr int A_SLICE_BYTES = TC_M * TC_K;  // 16 KB
constexpr int B_SLICE_BYTES = TC_N * TC_K;  // 8 KB

// ---------------------------------------------------------------------------------------------
// slicing kernels
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int exponent_of(double amax) {
  // amax = f * 2^e with f in [0.5, 1)  ->  |x| < 2^e for every |x| <= amax
  if (amax == 0.0) return 0;
  const int hi = __double2hiint(amax);
  return ((hi >> 20) & 0x7ff) - 1022;
}
__device__ __forceinline__ double pow2(int k) {  // 2^k for k in the normal range
  return __hiloint2double((1023 + k) << 20, 0);
}

// one warp per row: L[rows][K] (ld = ldl) -> Lq[s][kb][row][128] int8, scale[row] = 2^(e-6)
__global__ void __launch_bounds__(256) k_slice_rows(const double* __restrict__ L, int64_t ldl,
                                                    int64_t rows, int K, int S, int8_t* __restrict__ Lq,
                                                    int64_t rows_pad, double* __restrict__ scale) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int KB = (K + TC_K - 1) / TC_K;
  for (int64_t r = warp; r < rows; r += nwarps) {
    double amax = 0.0;
    for (int k = lane; k < K; k += 32) amax = fmax(amax, fabs(L[r * ldl + k]));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, off));
    const int e = exponent_of(amax);
    if (lane == 0) scale[r] = pow2(e - 6);
    const double up = pow2(6 - e);
    for (int k = lane; k < KB * TC_K; k += 32) {
      double t = (k < K) ? L[r * ldl + k] * up : 0.0;  // |t| < 64, exact
      const int kb = k / TC_K, kk = k % TC_K;
      for (int s = 0; s < S; ++s) {
        const double q = rint(t);
        Lq[(((int64_t)s * KB + kb) * rows_pad + r) * TC_K + kk] = (int8_t)(int)q;
        t = (t - q) * 128.0;  // exact: |t - q| <= 0.5
      }
    }
  }
}

// Rn[K][cols] (ld = ldr) -> Rq[s][kb][col][128] int8, scale[col] = 2^(e-6).
// One CTA per 32 columns: the K x 32 panel is read with coalesced 256-byte rows into shared memory,
// then ONE WARP PER COLUMN slices it - lane l owns k = 4l .. 4l+3 of the 128-deep k-block - so each
// (slice, column) is written as one 128-byte line.  (A thread-per-column version stored 16 bytes at
// a 128-byte stride: 8x the sectors, once per step over the whole pivot row.)
constexpr int SC_COLS = 32;
__global__ void __launch_bounds__(256) k_slice_cols(const double* __restrict__ Rn, int64_t ldr,
                                                    int64_t cols, int K, int S, int8_t* __restrict__ Rq,
                                                    int64_t cols_pad, double* __restrict__ scale,
                                                    const uint32_t* flags, uint32_t expect,
                                                    const JbCtl* ctl) {
  __shared__ double tile[TC_K][SC_COLS + 1];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t j0 = (int64_t)blockIdx.x * SC_COLS;
  if (ctl && (ctl->singular | ctl->timeout_flag)) return;
  if (flags) {  // wait until the pivot-row tile holding my columns has landed (owner's SCALE);
                // watchdog like every other cross-GPU wait (a dead peer must not hang this GPU)
    if (tid == 0)
      wait_flag_ge(flags + j0 / JB_BN, expect, ctl ? const_cast<int*>(&ctl->timeout_flag) : nullptr,
                   ctl ? ctl->spin_timeout_ns : 0);
    __syncthreads();
    if (ctl && *(volatile const int*)&ctl->timeout_flag) return;
  }
  const int KB = (K + TC_K - 1) / TC_K;
  for (int kb = 0; kb < KB; ++kb) {
    __syncthreads();
    for (int e = tid; e < TC_K * SC_COLS; e += 256) {
      const int k = e / SC_COLS, c = e % SC_COLS;
      const int kg = kb * TC_K + k;
      tile[k][c] = (kg < K && j0 + c < cols) ? Rn[(int64_t)kg * ldr + j0 + c] : 0.0;
    }
    __syncthreads();
    for (int c = warp; c < SC_COLS; c += 8) {
      const int64_t j = j0 + c;
      if (j >= cols) break;
      double t[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) t[i] = tile[4 * lane + i][c];
      // the column's scale comes from ALL k-blocks; KB > 1 callers pass one k-block per launch
      double amax = fmax(fmax(fabs(t[0]), fabs(t[1])), fmax(fabs(t[2]), fabs(t[3])));
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, off));
      const int ex = exponent_of(amax);
      if (lane == 0 && kb == 0) scale[j] = pow2(ex - 6);
      const double up = pow2(6 - ex);
#pragma unroll
      for (int i = 0; i < 4; ++i) t[i] *= up;
      for (int sl = 0; sl < S; ++sl) {
        uint32_t w = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const double q = rint(t[i]);
          w |= ((uint32_t)(uint8_t)(int8_t)(int)q) << (8 * i);
          t[i] = (t[i] - q) * 128.0;
        }
        reinterpret_cast<uint32_t*>(Rq + (((int64_t)sl * KB + kb) * cols_pad + j) * TC_K)[lane] = w;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// tcgen05 helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, int x, int y, int z,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
      : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, version 1)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);          // start address
  d |= (uint64_t)1 << 16;                              // leading byte offset (unused for K-major SW128)
  d |= (uint64_t)(1024 >> 4) << 32;                    // stride byte offset: 8 rows x 128 B
  d |= (uint64_t)1 << 46;                              // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                              // layout type SWIZZLE_128B
  return d;
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u), "r"(0u), "r"(0u),
      "r"(0u)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, int32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
                 "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
// exact int32 -> fp64 without I2F: bits(2^52 + (x + 2^31)) then subtract 2^52 + 2^31
__device__ __forceinline__ double i32_to_f64(int x) {
  return __hiloint2double(0x43300000, x ^ 0x80000000) - 4503601774854144.0;
}
__device__ __forceinline__ void tmem_wait_ld() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct Tc05Params {
  double* C;              // rows x cols fp64, ld = ldc (updated in place)
  int64_t ldc;
  const double* rscale;   // [rows]  2^(eL-6)
  const double* cscale;   // [cols]  2^(eR-6)
  int64_t rows, cols;     // rows: multiple of 128
  int KB, S;
  int rt_count, ct_count; // tiles
  // rows to skip (block row t / swapped-in row are handled by the fp64 kernel): [skip0, skip0+skiplen)
  int64_t skip0, skiplen, skip1, skiplen1;
  const JbCtl* ctl;       // optional: early-out when the matrix was found singular
  const JbStepInfo* stepinfo;  // optional (engine): swapped-in row comes from stepinfo[t]
  int t, m, rank;
  int col_t0, col_t1;     // column slot t: accumulate from zero instead of C (engine only; -1 = none)
  int dbg;                // timing experiments only: 1 = no C store, 2 = plain store instead of reduce-add
};

__global__ void __launch_bounds__(TC_THREADS, 1)
k_tc05_update(const Tc05Params P, const __grid_constant__ CUtensorMap tmA,
              const __grid_constant__ CUtensorMap tmB) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar_full, bar_smem_free, bar_tmem_full, bar_tmem_free;
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) double cs_tile[2][TC_N];  // column scales of the tile in flight (bulk-copied)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (P.ctl && (P.ctl->singular | P.ctl->timeout_flag)) return;

  // rows excluded from this kernel
  int64_t skipA0 = P.skip0, skipA1 = P.skip0 + P.skiplen, skipB0 = P.skip1, skipB1 = P.skip1 + P.skiplen1;
  if (P.stepinfo) {
    const JbStepInfo si = P.stepinfo[P.t];
    if (si.do_swap && si.piv_rank == P.rank) {
      skipB0 = (int64_t)si.piv_slot * P.m;
      skipB1 = skipB0 + P.m;
    }
  }

  unsigned char* sA = smem;                                   // S slices x 16 KB
  unsigned char* sB = smem + (size_t)P.S * A_SLICE_BYTES;     // S slices x 8 KB

  if (tid == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_smem_free, 1);
    mbar_init(&bar_tmem_full, 1);
    mbar_init(&bar_tmem_free, 4);  // one arrive per epilogue warp
    mbar_fence_init();
  }
  if (warp == 1) {  // TMEM: all 512 columns (S accumulators of 64 columns)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
        smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  const int total = P.rt_count * P.ct_count;
  auto row_skipped = [&](int rt) {
    const int64_t r0 = (int64_t)rt * TC_M;
    return (r0 >= skipA0 && r0 < skipA1) || (r0 >= skipB0 && r0 < skipB1);
  };

  if (warp == 0) {
    // ===================================================================== TMA producer
    if (lane == 0) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      uint32_t n = 0, pitem = 0;
      for (int idx = blockIdx.x; idx < total; idx += gridDim.x, ++pitem) {
        const int rt = idx % P.rt_count, ct = idx / P.rt_count;
        if (row_skipped(rt)) {
          --pitem;
          continue;
        }
        for (int kb = 0; kb < P.KB; ++kb, ++n) {
          mbar_wait(&bar_smem_free, (n & 1) ^ 1);
          mbar_expect_tx(&bar_full, (uint32_t)P.S * (A_SLICE_BYTES + B_SLICE_BYTES) +
                                        (kb == 0 ? (uint32_t)(TC_N * 8) : 0u));
          if (kb == 0) {
            asm volatile(
                "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                ::"r"(smem_u32(&cs_tile[pitem & 1][0])), "l"(P.cscale + (int64_t)ct * TC_N),
                "r"((uint32_t)(TC_N * 8)), "r"(smem_u32(&bar_full))
                : "memory");
          }
          for (int s = 0; s < P.S; ++s) {
            tma_load_3d(sA + (size_t)s * A_SLICE_BYTES, &tmA, 0, rt * TC_M, s * P.KB + kb, &bar_full);
            tma_load_3d(sB + (size_t)s * B_SLICE_BYTES, &tmB, 0, ct * TC_N, s * P.KB + kb, &bar_full);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      // instruction descriptor: D = S32, A = B = signed int8, both K-major, N = 64, M = 128
      const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_N >> 3) << 17) |
                             ((uint32_t)(TC_M >> 4) << 24);
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
      uint32_t n = 0, item = 0;
      for (int idx = blockIdx.x; idx < total; idx += gridDim.x) {
        const int rt = idx % P.rt_count;
        if (row_skipped(rt)) continue;
        mbar_wait(&bar_tmem_free, (item & 1) ^ 1);  // epilogue has drained the accumulators
        tc_fence_after();
        for (int kb = 0; kb < P.KB; ++kb, ++n) {
          mbar_wait(&bar_full, n & 1);
          tc_fence_after();
          for (int d = 0; d < P.S; ++d) {
            for (int s = 0; s <= d; ++s) {
              const int u = d - s;
              const uint64_t ad = umma_desc_sw128(a0 + s * A_SLICE_BYTES);
              const uint64_t bd = umma_desc_sw128(b0 + u * B_SLICE_BYTES);
#pragma unroll
              for (int k4 = 0; k4 < TC_K / 32; ++k4) {
                // advance 32 bytes (one UMMA_K of int8) inside the 128-byte swizzle atom
                umma_i8(tmem_base + d * TC_N, ad + (uint64_t)(k4 * 2), bd + (uint64_t)(k4 * 2), idesc,
                        (kb > 0 || s > 0 || k4 > 0) ? 1u : 0u);
              }
            }
          }
          umma_commit(&bar_smem_free);  // operands consumed -> producer may refill
        }
        umma_commit(&bar_tmem_full);    // all MMAs of the tile retired -> epilogue
        ++item;
      }
    }
  } else {
    // ===================================================================== epilogue warps 2..5
    const int q = warp & 3;                      // TMEM lane quarter this warp may access
    const int row_in_tile = q * 32 + lane;
    uint32_t item = 0;
    for (int idx = blockIdx.x; idx < total; idx += gridDim.x) {
      const int rt = idx % P.rt_count, ct = idx / P.rt_count;
      if (row_skipped(rt)) continue;
      const int64_t r = (int64_t)rt * TC_M + row_in_tile;
      const double rs = P.rscale[r];
      double* crow = P.C + r * P.ldc;
      const int64_t col0 = (int64_t)ct * TC_N;
      // prefetch my 64 C values (512 contiguous bytes of one row) while the MMAs of this tile run;
      // TMEM lanes are rows, so the thread <-> row mapping is fixed and every thread must hide its
      // own DRAM latency
      double2 cpre[TC_N / 2];
#pragma unroll
      for (int i = 0; i < TC_N / 2; ++i) {
        const int64_t c = col0 + 2 * i;
        double2 cv = make_double2(0.0, 0.0);
        if (c < P.cols) cv = *reinterpret_cast<const double2*>(crow + c);
        if (c >= P.col_t0 && c < P.col_t1) cv.x = 0.0;
        if (c + 1 >= P.col_t0 && c + 1 < P.col_t1) cv.y = 0.0;
        cpre[i] = cv;
      }
      mbar_wait(&bar_tmem_full, item & 1);
      tc_fence_after();
#pragma unroll
      for (int c8 = 0; c8 < TC_N; c8 += 8) {
        int32_t v[TC_SMAX][8];
#pragma unroll
        for (int d = 0; d < TC_SMAX; ++d) {
          if (d < P.S) {
            tmem_ld8(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(d * TC_N + c8), v[d]);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[d][i] = 0;
          }
        }
        tmem_wait_ld();
        const int64_t col = col0 + c8;
        if (col < P.cols) {
          double out[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            double acc = 0.0;
#pragma unroll
            for (int d = TC_SMAX - 1; d >= 0; --d) acc = fma(acc, 0.0078125, (double)v[d][i]);  // * 2^-7
            out[i] = acc;
          }
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const int64_t c = col + i;
            if (c < P.cols) {
              double2 cv = cpre[(c8 + i) / 2];
              cv.x = fma(out[i], rs * cs_tile[item & 1][c8 + i], cv.x);
              cv.y = fma(out[i + 1], rs * cs_tile[item & 1][c8 + i + 1], cv.y);
              *reinterpret_cast<double2*>(crow + c) = cv;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_tmem_free);
      ++item;
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

// ---------------------------------------------------------------------------------------------
// v2: A slices resident per row tile, 128 x 32 column tiles, double-buffered TMEM accumulators and
// B stages (MMA of tile i+1 overlaps the epilogue of tile i), and NO C traffic through the SM at
// all: the epilogue writes the scaled fp64 tile into a SWIZZLE_128B staging buffer and one thread
// issues a TMA reduce-add (cp.reduce.async.bulk.tensor ... .add, SASS UTMAREDG) that performs
// C += tile in L2.  K is exactly one 128-byte k-block per launch (m = 256 -> two launches).
// ---------------------------------------------------------------------------------------------
constexpr int V2_N = 32;
constexpr int V2_B_SLICE = V2_N * TC_K;              // 4 KB
constexpr int V2_B_STAGE = TC_SMAX * V2_B_SLICE;     // 32 KB
constexpr int V2_STG_HALF = TC_M * 16 * 8;           // 16 KB: 128 rows x 16 fp64 columns
constexpr int V2_SMEM = TC_SMAX * A_SLICE_BYTES + 2 * V2_B_STAGE + 2 * V2_STG_HALF;  // 224 KB

__device__ __forceinline__ int map_valid_rt(int vrt, int a0, int a1, int b0, int b1) {
  int rt = vrt;
  if (a1 > a0 && rt >= a0) rt += a1 - a0;
  if (b1 > b0 && rt >= b0) rt += b1 - b0;
  return rt;
}

template <int S_>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_tc05_update_v2(const Tc05Params P, const __grid_constant__ CUtensorMap tmA,
                 const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmC) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t a_full, a_free, b_full[2], b_free[2], t_full[2], t_free[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) double cs_tile[8][V2_N];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (P.ctl && (P.ctl->singular | P.ctl->timeout_flag)) return;

  // excluded row tiles (block row t; swapped-in row), as sorted, disjoint tile intervals
  int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
  if (P.skiplen > 0) {
    a0 = (int)(P.skip0 / TC_M);
    a1 = (int)((P.skip0 + P.skiplen) / TC_M);
  }
  if (P.stepinfo) {
    const JbStepInfo si = P.stepinfo[P.t];
    if (si.do_swap && si.piv_rank == P.rank) {
      b0 = (int)(((int64_t)si.piv_slot * P.m) / TC_M);
      b1 = (int)(((int64_t)(si.piv_slot + 1) * P.m) / TC_M);
    }
  }
  if (b1 > b0 && (a1 <= a0 || b0 < a0)) {  // order the intervals
    int t0 = a0, t1 = a1;
    a0 = b0; a1 = b1; b0 = t0; b1 = t1;
  }
  const int vrt_count = P.rt_count - (a1 - a0) - (b1 - b0);
  const int64_t total = (int64_t)vrt_count * P.ct_count;
  const int64_t lo = total * blockIdx.x / gridDim.x, hi = total * (blockIdx.x + 1) / gridDim.x;

  unsigned char* sA = smem;
  unsigned char* sB = smem + TC_SMAX * A_SLICE_BYTES;
  unsigned char* sStg = sB + 2 * V2_B_STAGE;

  if (tid == 0) {
    mbar_init(&a_full, 1);
    mbar_init(&a_free, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_free[i], 1);
      mbar_init(&t_full[i], 1);
      mbar_init(&t_free[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
        smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ===================================================================== TMA producer
    if (lane == 0 && lo < hi) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      int cur_rt = -1;
      uint32_t na = 0, nb = 0;
      for (int64_t idx = lo; idx < hi; ++idx, ++nb) {
        const int vrt = (int)(idx / P.ct_count), ct = (int)(idx % P.ct_count);
        const int rt = map_valid_rt(vrt, a0, a1, b0, b1);
        if (rt != cur_rt) {
          mbar_wait(&a_free, (na & 1) ^ 1);  // MMAs that read the previous A tile have retired
          mbar_expect_tx(&a_full, (uint32_t)P.S * A_SLICE_BYTES);
          for (int s = 0; s < P.S; ++s)
            tma_load_3d(sA + (size_t)s * A_SLICE_BYTES, &tmA, 0, rt * TC_M, s, &a_full);
          ++na;
          cur_rt = rt;
        }
        const int st = nb & 1;
        mbar_wait(&b_free[st], ((nb >> 1) & 1) ^ 1);
        mbar_expect_tx(&b_full[st], (uint32_t)P.S * V2_B_SLICE + (uint32_t)(V2_N * 8));
        for (int s = 0; s < P.S; ++s)
          tma_load_3d(sB + (size_t)st * V2_B_STAGE + (size_t)s * V2_B_SLICE, &tmB, 0, ct * V2_N, s,
                      &b_full[st]);
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
            ::"r"(smem_u32(&cs_tile[nb & 7][0])), "l"(P.cscale + (int64_t)ct * V2_N),
            "r"((uint32_t)(V2_N * 8)), "r"(smem_u32(&b_full[st]))
            : "memory");
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    if (lane == 0 && lo < hi) {
      const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(V2_N >> 3) << 17) |
                             ((uint32_t)(TC_M >> 4) << 24);
      const uint32_t abase = smem_u32(sA), bbase = smem_u32(sB);
      int cur_rt = -1;
      uint32_t na = 0, nb = 0;
      for (int64_t idx = lo; idx < hi; ++idx, ++nb) {
        const int rt = map_valid_rt((int)(idx / P.ct_count), a0, a1, b0, b1);
        if (rt != cur_rt) {
          mbar_wait(&a_full, na & 1);
          ++na;
          cur_rt = rt;
        }
        const int tb = nb & 1, st = nb & 1;
        mbar_wait(&t_free[tb], ((nb >> 1) & 1) ^ 1);  // epilogue drained this accumulator set
        mbar_wait(&b_full[st], (nb >> 1) & 1);
        tc_fence_after();
        // The issue loop runs in ONE thread, so it must be straight-line code: S is a template
        // parameter, all 36 x 4 descriptors are base + compile-time offsets (16-byte units), and the
        // accumulate flags are constants.  (A rolled loop cost ~75 cycles of integer work per MMA
        // against a 16-cycle MMA.)
        const uint64_t adesc0 = umma_desc_sw128(abase);
        const uint64_t bdesc0 = umma_desc_sw128(bbase + st * V2_B_STAGE);
        const uint32_t dbase = tmem_base + tb * (TC_SMAX * V2_N);
        if (P.dbg != 3)
        // One wide MMA per (slice s, k-step): A_s * [B_0 | ... | B_{S-1-s}] lands on the adjacent
        // accumulators D[s .. S-1] (see k_tc05_update_v3 for the reasoning).
#pragma unroll
        for (int k4 = 0; k4 < TC_K / 32; ++k4) {
#pragma unroll
          for (int s = 0; s < S_; ++s) {
            constexpr uint32_t kIdescBase = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_M >> 4) << 24);
            const uint32_t idesc_s = kIdescBase | ((uint32_t)(((S_ - s) * V2_N) >> 3) << 17);
            umma_i8(dbase + s * V2_N, adesc0 + (uint64_t)(s * (A_SLICE_BYTES / 16) + k4 * 2),
                    bdesc0 + (uint64_t)(k4 * 2), idesc_s, (s > 0 || k4 > 0) ? 1u : 0u);
          }
        }
        umma_commit(&b_free[st]);
        umma_commit(&t_full[tb]);
        const bool last_of_rt =
            (idx + 1 >= hi) || map_valid_rt((int)((idx + 1) / P.ct_count), a0, a1, b0, b1) != rt;
        if (last_of_rt) umma_commit(&a_free);
      }
    }
  } else {
    // ===================================================================== epilogue warps 2..5
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    const int ep_tid = tid - 64;
    uint32_t nb = 0;
    for (int64_t idx = lo; idx < hi; ++idx, ++nb) {
      const int vrt = (int)(idx / P.ct_count), ct = (int)(idx % P.ct_count);
      const int rt = map_valid_rt(vrt, a0, a1, b0, b1);
      const int tb = nb & 1;
      const double rs7 = P.rscale[(int64_t)rt * TC_M + row_in_tile] * 0.0078125;  // row scale * 2^-7
      mbar_wait(&t_full[tb], (nb >> 1) & 1);
      tc_fence_after();
      const double* cs = cs_tile[nb & 7];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        unsigned char* stg = sStg + h * V2_STG_HALF;
        // the reduce-store that read this staging half (two halves ago) must have finished reading
        if (ep_tid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        named_bar_sync(2, 128);
#pragma unroll
        for (int c8 = 0; c8 < 16; c8 += 8) {
          int32_t v[TC_SMAX][8];
#pragma unroll
          for (int d = 0; d < TC_SMAX; ++d) {
            if (d < P.S) {
              tmem_ld8(tmem_base + ((uint32_t)(q * 32) << 16) +
                           (uint32_t)(tb * (TC_SMAX * V2_N) + d * V2_N + h * 16 + c8), v[d]);
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) v[d][i] = 0;
            }
          }
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            // I2F.F64 runs at ~6-16 conversions/clk/SM on B200 (measured) and would dominate the
            // epilogue, so: (1) adjacent diagonals are merged in int32 (|D_d| < 2^22, so
            // D_2j * 128 + D_2j+1 fits), (2) int32 -> fp64 goes through the 2^52 magic constant
            // (one LOP3 + one exact DADD on the full-rate FP64 pipe).
            double o0 = 0.0, o1 = 0.0;
            if (P.dbg != 4)
#pragma unroll
            for (int j = TC_SMAX / 2 - 1; j >= 0; --j) {
              const int w0 = v[2 * j][i] * 128 + v[2 * j + 1][i];
              const int w1 = v[2 * j][i + 1] * 128 + v[2 * j + 1][i + 1];
              o0 = fma(o0, 6.103515625e-05, i32_to_f64(w0));  // * 2^-14
              o1 = fma(o1, 6.103515625e-05, i32_to_f64(w1));
            }
            o0 *= rs7 * cs[h * 16 + c8 + i];
            o1 *= rs7 * cs[h * 16 + c8 + i + 1];
            const int chunk = ((c8 + i) >> 1) ^ (row_in_tile & 7);  // SWIZZLE_128B
            *reinterpret_cast<double2*>(stg + row_in_tile * 128 + chunk * 16) = make_double2(o0, o1);
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        named_bar_sync(3, 128);
        if (ep_tid == 0 && P.dbg == 2) {
          asm volatile(
              "cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];"
              ::"l"(reinterpret_cast<uint64_t>(&tmC)), "r"(ct * V2_N + h * 16), "r"(rt * TC_M),
              "r"(smem_u32(stg))
              : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        } else if (ep_tid == 0 && P.dbg != 1) {
          asm volatile(
              "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
              ::"l"(reinterpret_cast<uint64_t>(&tmC)), "r"(ct * V2_N + h * 16), "r"(rt * TC_M),
              "r"(smem_u32(stg))
              : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&t_free[tb]);
    }
    if (ep_tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

// ---------------------------------------------------------------------------------------------
// v3 (default for S = 8; JB_TC05_NO_V3=1 falls back to v2): A slices live in TENSOR MEMORY.  In the SS form every
// tcgen05.mma re-reads its 4 KB A operand from shared memory, which at the small N that TMEM
// capacity allows costs ~5x the math (profiles/README.md).  Here the epilogue threads (thread =
// row = TMEM lane) load their rows of the S slices from L2 once per row tile and tcgen05.st them
// into TMEM columns [0, 256); the MMAs use the TS form (A from TMEM, B from smem through a
// 4-stage TMA ring); the 8 accumulators of one 128 x 32 tile occupy columns [256, 512).
// ---------------------------------------------------------------------------------------------
constexpr int V3_STAGES = 4;
constexpr int V3_SMEM = V3_STAGES * V2_B_STAGE + 2 * V2_STG_HALF;  // 160 KB
constexpr int V3_ACC0 = 256;                                        // first accumulator column

__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, {%5, %6, %7, %8}, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u), "r"(0u), "r"(0u),
      "r"(0u)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
      "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
      "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
      "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}

template <int S_>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_tc05_update_v3(const Tc05Params P, const int8_t* __restrict__ Lq, int64_t rows_pad,
                 const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmC) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t a_full, a_free, b_full[V3_STAGES], b_free[V3_STAGES], t_full, t_free;
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) double cs_tile[8][V2_N];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (P.ctl && (P.ctl->singular | P.ctl->timeout_flag)) return;

  int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
  if (P.skiplen > 0) {
    a0 = (int)(P.skip0 / TC_M);
    a1 = (int)((P.skip0 + P.skiplen) / TC_M);
  }
  if (P.stepinfo) {
    const JbStepInfo si = P.stepinfo[P.t];
    if (si.do_swap && si.piv_rank == P.rank) {
      b0 = (int)(((int64_t)si.piv_slot * P.m) / TC_M);
      b1 = (int)(((int64_t)(si.piv_slot + 1) * P.m) / TC_M);
    }
  }
  if (b1 > b0 && (a1 <= a0 || b0 < a0)) {
    int t0 = a0, t1 = a1;
    a0 = b0; a1 = b1; b0 = t0; b1 = t1;
  }
  const int vrt_count = P.rt_count - (a1 - a0) - (b1 - b0);
  const int64_t total = (int64_t)vrt_count * P.ct_count;
  const int64_t lo = total * blockIdx.x / gridDim.x, hi = total * (blockIdx.x + 1) / gridDim.x;

  unsigned char* sB = smem;
  unsigned char* sStg = smem + V3_STAGES * V2_B_STAGE;

  if (tid == 0) {
    mbar_init(&a_full, 4);   // one arrive per epilogue warp after its tcgen05.st are done
    mbar_init(&a_free, 1);
    for (int i = 0; i < V3_STAGES; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_free[i], 1);
    }
    mbar_init(&t_full, 1);
    mbar_init(&t_free, 4);
    mbar_fence_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
        smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ===================================================================== TMA producer (B only)
    if (lane == 0 && lo < hi) {
      tma_prefetch_desc(&tmB);
      uint32_t nb = 0;
      for (int64_t idx = lo; idx < hi; ++idx, ++nb) {
        const int ct = (int)(idx % P.ct_count);
        const int st = nb % V3_STAGES;
        mbar_wait(&b_free[st], ((nb / V3_STAGES) & 1) ^ 1);
        mbar_expect_tx(&b_full[st], (uint32_t)S_ * V2_B_SLICE + (uint32_t)(V2_N * 8));
        for (int s = 0; s < S_; ++s)
          tma_load_3d(sB + (size_t)st * V2_B_STAGE + (size_t)s * V2_B_SLICE, &tmB, 0, ct * V2_N, s,
                      &b_full[st]);
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
            ::"r"(smem_u32(&cs_tile[nb & 7][0])), "l"(P.cscale + (int64_t)ct * V2_N),
            "r"((uint32_t)(V2_N * 8)), "r"(smem_u32(&b_full[st]))
            : "memory");
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer (TS form)
    if (lane == 0 && lo < hi) {
      const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(V2_N >> 3) << 17) |
                             ((uint32_t)(TC_M >> 4) << 24);
      const uint32_t bbase = smem_u32(sB);
      int cur_rt = -1;
      uint32_t na = 0, nb = 0;
      for (int64_t idx = lo; idx < hi; ++idx, ++nb) {
        const int rt = map_valid_rt((int)(idx / P.ct_count), a0, a1, b0, b1);
        if (rt != cur_rt) {
          mbar_wait(&a_full, na & 1);  // the epilogue warps have stored this row tile's A slices
          ++na;
          cur_rt = rt;
        }
        const int st = nb % V3_STAGES;
        mbar_wait(&t_free, (nb & 1) ^ 1);
        mbar_wait(&b_full[st], (nb / V3_STAGES) & 1);
        tc_fence_after();
        const uint64_t bdesc0 = umma_desc_sw128(bbase + st * V2_B_STAGE);
        const uint32_t dbase = tmem_base + V3_ACC0;
        // The cost of a tcgen05.mma is dominated by delivering its 128 x 32-byte A operand (~64-80
        // cycles, independent of N), so N must be large.  The accumulators D[d] of consecutive
        // diagonals are adjacent TMEM column blocks and the B slices are adjacent 32-row blocks of
        // one K-major smem matrix, hence  A_s * [B_0 | B_1 | ... | B_{S-1-s}]  is ONE MMA with
        // N = 32 (S - s) that lands exactly on D[s .. S-1]:  4 S instructions per tile instead of
        // 2 S (S + 1).
#pragma unroll
        for (int k4 = 0; k4 < TC_K / 32; ++k4) {
#pragma unroll
          for (int s = 0; s < S_; ++s) {
            constexpr uint32_t kIdescBase = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_M >> 4) << 24);
            const uint32_t idesc_s = kIdescBase | ((uint32_t)(((S_ - s) * V2_N) >> 3) << 17);
            umma_i8_ts(dbase + s * V2_N, tmem_base + s * 32 + k4 * 8, bdesc0 + (uint64_t)(k4 * 2), idesc_s,
                       (s > 0 || k4 > 0) ? 1u : 0u);
          }
        }
        umma_commit(&b_free[st]);
        umma_commit(&t_full);
        const bool last_of_rt =
            (idx + 1 >= hi) || map_valid_rt((int)((idx + 1) / P.ct_count), a0, a1, b0, b1) != rt;
        if (last_of_rt) umma_commit(&a_free);
      }
    }
  } else {
    // ===================================================================== epilogue warps 2..5
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    const int ep_tid = tid - 64;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    int cur_rt = -1;
    uint32_t na = 0, nb = 0;
    double rs7 = 0.0;
    for (int64_t idx = lo; idx < hi; ++idx, ++nb) {
      const int vrt = (int)(idx / P.ct_count), ct = (int)(idx % P.ct_count);
      const int rt = map_valid_rt(vrt, a0, a1, b0, b1);
      if (rt != cur_rt) {
        // (re)load the A slices of this row tile into TMEM: thread = row = TMEM lane
        mbar_wait(&a_free, (na & 1) ^ 1);  // MMAs reading the previous A have retired
        tc_fence_after();
        const int64_t r = (int64_t)rt * TC_M + row_in_tile;
#pragma unroll 1
        for (int s = 0; s < S_; ++s) {
          const uint4* src = reinterpret_cast<const uint4*>(Lq + ((int64_t)s * rows_pad + r) * TC_K);
          uint32_t v[32];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint4 x = src[i];
            v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
          }
          tmem_st32(tmem_base + lane_base + (uint32_t)(s * 32), v);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full);
        rs7 = P.rscale[r] * 0.0078125;
        ++na;
        cur_rt = rt;
      }
      mbar_wait(&t_full, nb & 1);
      tc_fence_after();
      const double* cs = cs_tile[nb & 7];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        unsigned char* stg = sStg + h * V2_STG_HALF;
        if (ep_tid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        named_bar_sync(2, 128);
#pragma unroll
        for (int c8 = 0; c8 < 16; c8 += 8) {
          int32_t v[TC_SMAX][8];
#pragma unroll
          for (int d = 0; d < TC_SMAX; ++d) {
            if (d < S_) {
              tmem_ld8(tmem_base + lane_base + (uint32_t)(V3_ACC0 + d * V2_N + h * 16 + c8), v[d]);
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) v[d][i] = 0;
            }
          }
          tmem_wait_ld();
          if (h == 1 && c8 == 8) {  // last TMEM read of this tile: the MMA of the next tile may start
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&t_free);
          }
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            double o0 = 0.0, o1 = 0.0;
#pragma unroll
            for (int j = TC_SMAX / 2 - 1; j >= 0; --j) {
              const int w0 = v[2 * j][i] * 128 + v[2 * j + 1][i];
              const int w1 = v[2 * j][i + 1] * 128 + v[2 * j + 1][i + 1];
              o0 = fma(o0, 6.103515625e-05, i32_to_f64(w0));
              o1 = fma(o1, 6.103515625e-05, i32_to_f64(w1));
            }
            o0 *= rs7 * cs[h * 16 + c8 + i];
            o1 *= rs7 * cs[h * 16 + c8 + i + 1];
            const int chunk = ((c8 + i) >> 1) ^ (row_in_tile & 7);  // SWIZZLE_128B
            *reinterpret_cast<double2*>(stg + row_in_tile * 128 + chunk * 16) = make_double2(o0, o1);
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        named_bar_sync(3, 128);
        if (ep_tid == 0) {
          asm volatile(
              "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
              ::"l"(reinterpret_cast<uint64_t>(&tmC)), "r"(ct * V2_N + h * 16), "r"(rt * TC_M),
              "r"(smem_u32(stg))
              : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (ep_tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

// W(r, col0 .. col0+ncols) = 0 for every local row outside [skip0, skip0+skiplen): in the in-place
// elimination column slot t starts from the identity column, and the v2 kernel only ever adds.
__global__ void __launch_bounds__(256) k_zero_cols(double* W, int64_t ldw, int64_t rows, int64_t col0,
                                                   int ncols, int64_t skip0, int64_t skiplen,
                                                   const JbCtl* ctl) {
  if (ctl && (ctl->singular | ctl->timeout_flag)) return;
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    if (r >= skip0 && r < skip0 + skiplen) continue;
    for (int c = threadIdx.x; c < ncols; c += blockDim.x) W[r * ldw + col0 + c] = 0.0;
  }
}

int make_map_f64_sw128(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t ld);

int make_map_i8(CUtensorMap* map, const void* base, uint64_t rows, uint64_t planes, uint32_t box_rows) {
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      jb_set_error("cuTensorMapEncodeTiled entry point not available");
      return JB_ERR_CUDA;
    }
    fn = (EncodeTiledFn)p;
  }
  cuuint64_t gdim[3] = {(cuuint64_t)TC_K, rows, planes};
  cuuint64_t gstr[2] = {(cuuint64_t)TC_K, (cuuint64_t)TC_K * rows};
  cuuint32_t box[3] = {(cuuint32_t)TC_K, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    jb_set_error("cuTensorMapEncodeTiled(int8) failed: %d", (int)r);
    return JB_ERR_CUDA;
  }
  return JB_OK;
}

int make_map_f64_sw128(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t ld) {
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess) {
    jb_set_error("cuTensorMapEncodeTiled entry point not available");
    return JB_ERR_CUDA;
  }
  EncodeTiledFn fn = (EncodeTiledFn)p;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstr[1] = {ld * 8};
  cuuint32_t box[2] = {16, (cuuint32_t)TC_M};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    jb_set_error("cuTensorMapEncodeTiled(fp64 C) failed: %d", (int)r);
    return JB_ERR_CUDA;
  }
  return JB_OK;
}

}  // namespace

namespace {
}
size_t jb_tc05_smem_bytes(int S) { return (size_t)S * (A_SLICE_BYTES + B_SLICE_BYTES); }

// Workspace sizes for `rows` x K (left) and K x `cols` (right) operands.
size_t jb_tc05_lq_bytes(int64_t rows_pad, int K, int S) {
  return (size_t)S * ((K + TC_K - 1) / TC_K) * rows_pad * TC_K;
}
size_t jb_tc05_rq_bytes(int64_t cols_pad, int K, int S) {
  return (size_t)S * ((K + TC_K - 1) / TC_K) * cols_pad * TC_K;
}

cudaError_t jb_tc05_slice_rows(const double* L, int64_t ldl, int64_t rows, int K, int S, int8_t* Lq,
                               int64_t rows_pad, double* scale, cudaStream_t st) {
  if (rows == 0) return cudaSuccess;
  int grid = (int)std::min<int64_t>((rows + 7) / 8, 148 * 8);
  k_slice_rows<<<grid, 256, 0, st>>>(L, ldl, rows, K, S, Lq, rows_pad, scale);
  return cudaGetLastError();
}

cudaError_t jb_tc05_slice_cols(const double* Rn, int64_t ldr, int64_t cols, int K, int S, int8_t* Rq,
                               int64_t cols_pad, double* scale, const uint32_t* flags, uint32_t expect,
                               const JbCtl* ctl, cudaStream_t st) {
  if (cols == 0) return cudaSuccess;
  if (K > TC_K) return cudaErrorInvalidValue;  // one 128-deep k-block per call (the scale is per block)
  int grid = (int)((cols + SC_COLS - 1) / SC_COLS);
  k_slice_cols<<<grid, 256, 0, st>>>(Rn, ldr, cols, K, S, Rq, cols_pad, scale, flags, expect, ctl);
  return cudaGetLastError();
}

// One K <= 128 block: C += Lq * Rq with the scales of that block.  version 2 (default): resident-A,
// TMA reduce-add epilogue; the caller has zeroed column slot t beforehand (k_zero_cols).
// The three tensor maps of a launch depend only on the operand buffers and shapes, which an engine
// re-uses step after step: encode them once and look them up afterwards (cuTensorMapEncodeTiled is
// a driver call of several microseconds, three per launch, on a launch-bound path).
struct Tc05MapKey {
  const void *Lq, *Rq, *C;
  int64_t rows_pad, cols_pad, rows, cols, ldc;
  int S, dev, v1;
  bool operator==(const Tc05MapKey& o) const { return memcmp(this, &o, sizeof(*this)) == 0; }
};
struct Tc05MapEntry {
  Tc05MapKey key;
  CUtensorMap tmA, tmB, tmC;
};

int jb_tc05_launch(const Tc05Launch& a, cudaStream_t st) {
  static const bool use_v1 = getenv("JB_TC05_V1") != nullptr;
  static std::mutex mu;
  static std::vector<Tc05MapEntry> cache;
  int cur_dev = 0;
  cudaGetDevice(&cur_dev);
  Tc05MapKey key;
  memset(&key, 0, sizeof(key));
  key.Lq = a.Lq; key.Rq = a.Rq; key.C = a.C; key.rows_pad = a.rows_pad; key.cols_pad = a.cols_pad;
  key.rows = a.rows; key.cols = a.cols; key.ldc = a.ldc; key.S = a.S; key.dev = cur_dev; key.v1 = use_v1;
  CUtensorMap tmA, tmB, tmC;
  bool hit = false;
  {
    std::lock_guard<std::mutex> lock(mu);
    for (const Tc05MapEntry& en : cache)
      if (en.key == key) {
        tmA = en.tmA; tmB = en.tmB; tmC = en.tmC;
        hit = true;
        break;
      }
  }
  int rc = JB_OK;
  if (!hit) {
    rc = make_map_i8(&tmA, a.Lq, (uint64_t)a.rows_pad, (uint64_t)a.S, TC_M);
    if (rc) return rc;
    rc = make_map_i8(&tmB, a.Rq, (uint64_t)a.cols_pad, (uint64_t)a.S, use_v1 ? TC_N : V2_N);
    if (rc) return rc;
    memset(&tmC, 0, sizeof(tmC));
    if (!use_v1 && a.rows >= TC_M && a.cols > 0) {
      rc = make_map_f64_sw128(&tmC, a.C, (uint64_t)a.cols, (uint64_t)a.rows, (uint64_t)a.ldc);
      if (rc) return rc;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (cache.size() >= 64) cache.erase(cache.begin());
    Tc05MapEntry en;
    en.key = key; en.tmA = tmA; en.tmB = tmB; en.tmC = tmC;
    cache.push_back(en);
  }
  Tc05Params P;
  memset(&P, 0, sizeof(P));
  P.C = a.C; P.ldc = a.ldc; P.rscale = a.rscale; P.cscale = a.cscale; P.rows = a.rows; P.cols = a.cols;
  P.KB = 1; P.S = a.S; P.rt_count = (int)(a.rows / TC_M);
  P.skip0 = a.skip0; P.skiplen = a.skiplen; P.skip1 = 0; P.skiplen1 = 0;
  P.ctl = a.ctl; P.stepinfo = a.stepinfo; P.t = a.t; P.m = a.m; P.rank = a.rank;
  P.col_t0 = a.col_t0; P.col_t1 = a.col_t1;
  if (const char* d = getenv("JB_TC05_DBG")) P.dbg = atoi(d);
  static bool configured[JB_MAX_RANKS] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev % JB_MAX_RANKS]) {
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)jb_tc05_smem_bytes(TC_SMAX)));
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update_v2<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, V2_SMEM));
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update_v2<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, V2_SMEM));
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update_v2<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, V2_SMEM));
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update_v3<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, V3_SMEM));
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update_v3<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, V3_SMEM));
    JB_CHECK(cudaFuncSetAttribute(k_tc05_update_v3<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, V3_SMEM));
    configured[dev % JB_MAX_RANKS] = true;
  }
  if (P.rt_count == 0 || a.cols == 0) return JB_OK;
  if (use_v1) {
    P.ct_count = (int)((a.cols + TC_N - 1) / TC_N);
    k_tc05_update<<<a.grid, TC_THREADS, jb_tc05_smem_bytes(a.S), st>>>(P, tmA, tmB);
  } else {
    P.ct_count = (int)((a.cols + V2_N - 1) / V2_N);
    // column slot t: the v2 epilogue only adds, so the slot is cleared first
    if (a.col_t1 > a.col_t0 && a.zero_slot) {
      int grid = (int)std::min<int64_t>(a.rows, 148 * 8);
      k_zero_cols<<<grid, 256, 0, st>>>(a.C, a.ldc, a.rows, a.col_t0, a.col_t1 - a.col_t0, a.skip0,
                                       a.skiplen, a.ctl);
      JB_CHECK(cudaGetLastError());
    }
    static const bool use_v3 = getenv("JB_TC05_NO_V3") == nullptr;  // A-in-TMEM kernel (default)
    // S = 8 is the validated default; 7 / 6 slices (JB_TC05_SLICES) trade accuracy for speed and use the
    // v2 kernel unless JB_TC05_V3_ALL=1 (same code, other loop bounds; not measured this round)
    static const bool v3_all = getenv("JB_TC05_V3_ALL") != nullptr;
    if (a.S == 8 && use_v3) k_tc05_update_v3<8><<<a.grid, TC_THREADS, V3_SMEM, st>>>(P, a.Lq, a.rows_pad, tmB, tmC);
    else if (a.S == 7 && use_v3 && v3_all) k_tc05_update_v3<7><<<a.grid, TC_THREADS, V3_SMEM, st>>>(P, a.Lq, a.rows_pad, tmB, tmC);
    else if (a.S == 6 && use_v3 && v3_all) k_tc05_update_v3<6><<<a.grid, TC_THREADS, V3_SMEM, st>>>(P, a.Lq, a.rows_pad, tmB, tmC);
    else if (a.S == 8) k_tc05_update_v2<8><<<a.grid, TC_THREADS, V2_SMEM, st>>>(P, tmA, tmB, tmC);
    else if (a.S == 7) k_tc05_update_v2<7><<<a.grid, TC_THREADS, V2_SMEM, st>>>(P, tmA, tmB, tmC);
    else if (a.S == 6) k_tc05_update_v2<6><<<a.grid, TC_THREADS, V2_SMEM, st>>>(P, tmA, tmB, tmC);
    else {
      jb_set_error("tcgen05 update v2 is built for 6, 7 or 8 slices (got %d); set JB_TC05_V1=1 for others", a.S);
      return JB_ERR_ARG;
    }
  }
  JB_CHECK(cudaGetLastError());
  return JB_OK;
}

// Stand-alone entry for tests / tools: C (rows x cols) += L (rows x K) * Rn (K x cols), all fp64 in
// device memory, computed on the tcgen05 int8 path with S slices.  rows must be a multiple of 128.
extern "C" int jb_tc05_gemm_update(double* C, int64_t ldc, int64_t rows, int64_t cols, const double* L,
                                   int64_t ldl, const double* Rn, int64_t ldr, int32_t K, int32_t S,
                                   void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (rows % TC_M != 0 || S < 1 || S > TC_SMAX || K < 1 || (ldc & 1) || cols < 1 ||
      (reinterpret_cast<uintptr_t>(C) & 15)) {
    jb_set_error("jb_tc05_gemm_update: rows must be a multiple of 128, 1 <= S <= 8, even ldc");
    return JB_ERR_ARG;
  }
  const int64_t cols_pad = (cols + TC_N - 1) / TC_N * TC_N;
  int8_t *Lq = nullptr, *Rq = nullptr;
  double *rs = nullptr, *cs = nullptr;
  JB_CHECK(cudaMalloc((void**)&Lq, jb_tc05_lq_bytes(rows, TC_K, S)));
  JB_CHECK(cudaMalloc((void**)&Rq, jb_tc05_rq_bytes(cols_pad, TC_K, S)));
  JB_CHECK(cudaMalloc((void**)&rs, (size_t)rows * 8));
  JB_CHECK(cudaMalloc((void**)&cs, (size_t)cols_pad * 8));
  int rc = JB_OK;
  for (int k0 = 0; k0 < K && rc == JB_OK; k0 += TC_K) {  // one 128-wide k-block at a time
    const int kb = std::min(TC_K, K - k0);
    JB_CHECK(cudaMemsetAsync(Rq, 0, jb_tc05_rq_bytes(cols_pad, TC_K, S), st));
    JB_CHECK(cudaMemsetAsync(cs, 0, (size_t)cols_pad * 8, st));
    JB_CHECK(jb_tc05_slice_rows(L + k0, ldl, rows, kb, S, Lq, rows, rs, st));
    JB_CHECK(jb_tc05_slice_cols(Rn + (int64_t)k0 * ldr, ldr, cols, kb, S, Rq, cols_pad, cs, nullptr, 0, nullptr,
                                st));
    Tc05Launch a;
    memset(&a, 0, sizeof(a));
    a.C = C; a.ldc = ldc; a.rows = rows; a.cols = cols; a.Lq = Lq; a.Rq = Rq; a.rows_pad = rows;
    a.cols_pad = cols_pad; a.rscale = rs; a.cscale = cs; a.K = kb; a.S = S; a.grid = 148;
    a.col_t0 = -1; a.col_t1 = -1;
    rc = jb_tc05_launch(a, st);
  }
  cudaError_t e = cudaStreamSynchronize(st);
  cudaFree(Lq); cudaFree(Rq); cudaFree(rs); cudaFree(cs);
  if (rc) return rc;
  JB_CHECK(e);
  return JB_OK;
}
