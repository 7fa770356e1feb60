// K5 `gen_fill_norm`: generator fill of the block-cyclic local rows fused with the ||A||inf
// row-sum reduction, plus the small elementwise helpers (column un-permutation, max-reduce,
// peer gather of an X column panel).
//
// Reference: f / f_i (main.cpp:47-64), init_matrix (128-149), local_to_global (118-123),
// norm (643-667).  Rows/columns beyond n (ragged last block) are filled with the
// identity so the padded matrix is diag(A, I): its inverse is diag(A^-1, I) and no kernel needs a
// ragged-edge special case.  A padded block row can never win a pivot search before the last step
// (its candidate block has zero rows -> singular), which is the rule of main.cpp:967-970.
#include "common.cuh"
#include "engine_types.h"
#include "kernels.h"

namespace {

__device__ __forceinline__ void atomic_max_nonneg(double* addr, double v) {
  // non-negative IEEE doubles order like their bit patterns
  atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}

__device__ __forceinline__ double block_sum(double s) {
  __shared__ double red[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  double t = 0.0;
  if (warp == 0) {
    t = (lane < (blockDim.x >> 5)) ? red[lane] : 0.0;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
  }
  __syncthreads();
  return t;  // valid in warp 0
}

__global__ void __launch_bounds__(256) k_fill(const FillParams P) {
  const int64_t rows = (int64_t)P.rloc * P.m;
  for (int64_t lr = blockIdx.x; lr < rows; lr += gridDim.x) {
    const int64_t gi = ((lr / P.m) * P.p + P.rank) * P.m + lr % P.m;  // main.cpp:121
    double* row = P.W + lr * P.ldw;
    double s = 0.0;
    for (int64_t j = threadIdx.x; j < P.ldw; j += blockDim.x) {
      double v;
      if (gi < P.n && j < P.n) {
        v = jb_formula_eval(P.formula, gi, j, P.n);
        s += fabs(v);
      } else {
        v = (gi == j) ? 1.0 : 0.0;
      }
      row[j] = v;
    }
    s = block_sum(s);
    if (threadIdx.x == 0 && P.rowsum_max) atomic_max_nonneg(P.rowsum_max, s);
  }
}

__global__ void __launch_bounds__(256) k_rowsum_max(const double* W, int64_t rows, int64_t cols,
                                                    int64_t ld, double* out) {
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    double s = 0.0;
    for (int64_t j = threadIdx.x; j < cols; j += blockDim.x) s += fabs(W[r * ld + j]);
    s = block_sum(s);
    if (threadIdx.x == 0) atomic_max_nonneg(out, s);
  }
}

__global__ void __launch_bounds__(256) k_max_reduce(const double* v, int64_t count, double* out) {
  double mx = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * blockDim.x)
    mx = fmax(mx, fabs(v[i]));
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  if ((threadIdx.x & 31) == 0) atomic_max_nonneg(out, mx);
}

// In-place Gauss-Jordan leaves column pi(t) of the inverse in column slot t.  Undo the block-row
// interchanges on the columns, last swap first (row-local, one CTA per row).
__global__ void __launch_bounds__(128) k_unpermute(double* W, int64_t rows, int64_t ldw, int m,
                                                   int Nr, const JbStepInfo* si) {
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    double* row = W + r * ldw;
    for (int t = Nr - 1; t >= 0; --t) {
      const int pg = si[t].piv_g;
      if (pg == t || pg < 0) continue;  // uniform over the CTA
      for (int j = threadIdx.x; j < m; j += blockDim.x) {
        double a = row[(int64_t)t * m + j], b = row[(int64_t)pg * m + j];
        row[(int64_t)t * m + j] = b;
        row[(int64_t)pg * m + j] = a;
      }
      __syncthreads();
    }
  }
}

}  // namespace

cudaError_t jb_launch_fill(const FillParams& P, cudaStream_t st) {
  int64_t rows = (int64_t)P.rloc * P.m;
  if (rows == 0) return cudaSuccess;
  int grid = (int)(rows < 148 * 16 ? rows : 148 * 16);
  k_fill<<<grid, 256, 0, st>>>(P);
  return cudaGetLastError();
}

cudaError_t jb_launch_rowsum_max(const double* W, int64_t rows, int64_t cols, int64_t ld,
                                 double* out_max, cudaStream_t st) {
  if (rows == 0) return cudaSuccess;
  int grid = (int)(rows < 148 * 16 ? rows : 148 * 16);
  k_rowsum_max<<<grid, 256, 0, st>>>(W, rows, cols, ld, out_max);
  return cudaGetLastError();
}

cudaError_t jb_launch_max_reduce(const double* v, int64_t count, double* out, cudaStream_t st) {
  if (count == 0) return cudaSuccess;
  int64_t blocks = (count + 255) / 256;
  int grid = (int)(blocks < 148 * 8 ? blocks : 148 * 8);
  k_max_reduce<<<grid, 256, 0, st>>>(v, count, out);
  return cudaGetLastError();
}

cudaError_t jb_launch_unpermute(double* W, int64_t rows, int64_t ldw, int m, int Nr,
                                const JbStepInfo* stepinfo, cudaStream_t st) {
  if (rows == 0) return cudaSuccess;
  int grid = (int)(rows < 148 * 16 ? rows : 148 * 16);
  k_unpermute<<<grid, 128, 0, st>>>(W, rows, ldw, m, Nr, stepinfo);
  return cudaGetLastError();
}

// dst[b][i][j] = src[b][i][j] for i, j < m: blocks stored with pitch Kp (m rounded up to 16) are
// packed to pitch m in ONE launch (the stand-alone block-inversion entry points return m x m blocks)
__global__ void k_compact_blocks(const double* __restrict__ src, int Kp, double* __restrict__ dst, int m,
                                 int64_t nblocks) {
  const int64_t per = (int64_t)m * m;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nblocks * per;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = e / per;
    const int r = (int)((e - b * per) / m), c = (int)(e - b * per - (int64_t)r * m);
    dst[e] = src[(b * Kp + r) * Kp + c];
  }
}

cudaError_t jb_launch_compact_blocks(const double* src, int Kp, double* dst, int m, int64_t nblocks,
                                     cudaStream_t st) {
  if (nblocks <= 0) return cudaSuccess;
  const int64_t total = nblocks * m * m;
  const int grid = (int)(total / 256 + 1 < 148 * 8 ? total / 256 + 1 : 148 * 8);
  k_compact_blocks<<<grid, 256, 0, st>>>(src, Kp, dst, m, nblocks);
  return cudaGetLastError();
}

