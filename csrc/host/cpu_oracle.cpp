// CPU oracle: block Gauss-Jordan inversion with block-row pivoting, "intended semantics" of the
// reference (defect D1 fixed: the block row that WON the pivot search is the one
// that is scaled, broadcast and swapped).  Written independently of the GPU engine: it works on the
// explicit augmented system [A | X] like the reference does (3 n^3 flops), whereas the engine runs
// the in-place 2 n^3 form - agreement between the two is a real check.
// The p ranks of the block-cyclic distribution are emulated inside one process so that the
// rank-dependent tie rules of the MINPIV reduction (main.cpp:729-744, :1024) are reproduced.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../include/jordan_b200.h"

#define JB_EPS 1e-15

extern "C" {

// Gauss-Jordan with partial pivoting on [a | b], b := I first (inverse_block, main.cpp:746-820).
int jb_cpu_inverse_block(double* a, double* b, double norm, int64_t n) {
  for (int64_t i = 0; i < n * n; ++i) b[i] = 0.0;
  for (int64_t i = 0; i < n; ++i) b[i * n + i] = 1.0;
  for (int64_t k = 0; k < n; ++k) {
    int64_t piv = -1;
    double best = -1.0;
    for (int64_t i = k; i < n; ++i) {
      double v = fabs(a[i * n + k]);
      if (v > best) {  // strict: first maximum wins
        best = v;
        piv = i;
      }
    }
    if (piv != k && piv >= 0) {
      for (int64_t j = 0; j < n; ++j) {
        double t = a[k * n + j]; a[k * n + j] = a[piv * n + j]; a[piv * n + j] = t;
        t = b[k * n + j]; b[k * n + j] = b[piv * n + j]; b[piv * n + j] = t;
      }
    }
    if (fabs(a[k * n + k]) < JB_EPS * norm || fabs(norm) < JB_EPS) return 1;
    const double d = 1.0 / a[k * n + k];
    for (int64_t j = 0; j < n; ++j) {
      a[k * n + j] *= d;
      b[k * n + j] *= d;
    }
    for (int64_t i = 0; i < n; ++i) {
      if (i == k) continue;
      const double f = a[i * n + k];
      if (f == 0.0) continue;
      for (int64_t j = 0; j < n; ++j) {
        a[i * n + j] -= f * a[k * n + j];
        b[i * n + j] -= f * b[k * n + j];
      }
    }
  }
  return 0;
}

// MPI_Allreduce with the non-commutative user op MINPIV: r0 o r1 o ... o r(p-1), where
// (lo o hi) keeps hi unless lo is non-singular and strictly better.
int jb_cpu_minloc_fold(const double* norms, const int32_t* nonsing, int64_t p) {
  int64_t best = 0;
  for (int64_t r = 1; r < p; ++r) {
    bool lo_wins = nonsing[best] && (!nonsing[r] || norms[best] < norms[r]);
    if (!lo_wins) best = r;
  }
  return nonsing[best] ? (int)best : -1;
}

static double inf_norm(const double* a, int64_t rows, int64_t cols, int64_t ld) {
  double mx = 0.0;
  for (int64_t i = 0; i < rows; ++i) {
    double s = 0.0;
    for (int64_t j = 0; j < cols; ++j) s += fabs(a[i * ld + j]);
    if (s > mx) mx = s;
  }
  return mx;
}

int jb_cpu_block_jordan(double* a, double* x, int64_t n, int64_t m, int64_t p, int32_t* piv_out) {
  const int64_t Nr = jb_num_block_rows(n, m);
  const int64_t lh = n - (Nr - 1) * m;  // height of the ragged last block row
  const double norm_a = inf_norm(a, n, n, n);
  for (int64_t i = 0; i < n * n; ++i) x[i] = 0.0;
  for (int64_t i = 0; i < n; ++i) x[i * n + i] = 1.0;
  std::vector<double> E((size_t)m * m), F((size_t)m * m), Pinv((size_t)m * m);
  std::vector<double> prow_a((size_t)m * n), prow_x((size_t)m * n), tmp((size_t)m * n);
  std::vector<double> rk_norm((size_t)p);
  std::vector<int32_t> rk_ok((size_t)p), rk_row((size_t)p);

  auto bh = [&](int64_t I) { return I == Nr - 1 ? lh : m; };

  for (int64_t t = 0; t < Nr; ++t) {
    const int64_t w = bh(t);  // width of block column t
    // (1) per-rank search over candidates I >= t; the ragged row only at the last step
    for (int64_t r = 0; r < p; ++r) {
      rk_ok[r] = 0;
      rk_norm[r] = 0.0;
      rk_row[r] = -1;
    }
    for (int64_t I = t; I < Nr; ++I) {
      if (bh(I) != w) continue;  // ragged row vs full column (or vice versa): not square
      const int64_t r = I % p;
      for (int64_t i = 0; i < w; ++i)
        for (int64_t j = 0; j < w; ++j) E[i * w + j] = a[(I * m + i) * n + t * m + j];
      if (jb_cpu_inverse_block(E.data(), F.data(), norm_a, w)) continue;
      const double bn = inf_norm(F.data(), w, w, w);
      if (!rk_ok[r] || bn < rk_norm[r]) {  // strict <, first wins
        rk_ok[r] = 1;
        rk_norm[r] = bn;
        rk_row[r] = (int32_t)I;
      }
    }
    // (2) global min-loc with the reference's tie rule
    const int win = jb_cpu_minloc_fold(rk_norm.data(), rk_ok.data(), p);
    if (win < 0) return JB_ERR_SINGULAR;
    const int64_t P = rk_row[win];
    if (piv_out) piv_out[t] = (int32_t)P;
    // inverse of the winning block
    for (int64_t i = 0; i < w; ++i)
      for (int64_t j = 0; j < w; ++j) E[i * w + j] = a[(P * m + i) * n + t * m + j];
    jb_cpu_inverse_block(E.data(), Pinv.data(), norm_a, w);
    // (4) swap block rows t and P (both have height w)
    if (P != t) {
      for (int64_t i = 0; i < w; ++i) {
        double* ra = a + (t * m + i) * n; double* rb = a + (P * m + i) * n;
        memcpy(tmp.data(), ra, (size_t)n * 8); memcpy(ra, rb, (size_t)n * 8); memcpy(rb, tmp.data(), (size_t)n * 8);
        ra = x + (t * m + i) * n; rb = x + (P * m + i) * n;
        memcpy(tmp.data(), ra, (size_t)n * 8); memcpy(ra, rb, (size_t)n * 8); memcpy(rb, tmp.data(), (size_t)n * 8);
      }
    }
    // (5) pivot row <- Pinv * row, in both halves
    for (int64_t i = 0; i < w; ++i) {
      for (int64_t j = 0; j < n; ++j) {
        double sa = 0.0, sx = 0.0;
        for (int64_t k = 0; k < w; ++k) {
          sa += Pinv[i * w + k] * a[(t * m + k) * n + j];
          sx += Pinv[i * w + k] * x[(t * m + k) * n + j];
        }
        prow_a[i * n + j] = sa;
        prow_x[i * n + j] = sx;
      }
    }
    for (int64_t i = 0; i < w; ++i) {
      memcpy(a + (t * m + i) * n, prow_a.data() + i * n, (size_t)n * 8);
      memcpy(x + (t * m + i) * n, prow_x.data() + i * n, (size_t)n * 8);
    }
    // (6) eliminate block column t from every other row
    for (int64_t i = 0; i < n; ++i) {
      if (i >= t * m && i < t * m + w) continue;
      double* ra = a + i * n;
      double* rx = x + i * n;
      for (int64_t k = 0; k < w; ++k) {
        const double f = ra[t * m + k];
        if (f == 0.0) continue;
        const double* pa = prow_a.data() + k * n;
        const double* px = prow_x.data() + k * n;
        for (int64_t j = 0; j < n; ++j) {
          if (j < t * m || j >= t * m + w) ra[j] -= f * pa[j];
          rx[j] -= f * px[j];
        }
      }
      for (int64_t k = 0; k < w; ++k) ra[t * m + k] = 0.0;
    }
  }
  return JB_OK;
}

double jb_cpu_residual(const double* a, const double* x, int64_t n) {
  std::vector<double> row((size_t)n);
  double mx = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    for (int64_t j = 0; j < n; ++j) row[j] = 0.0;
    for (int64_t k = 0; k < n; ++k) {
      const double f = a[i * n + k];
      if (f == 0.0) continue;
      const double* xr = x + k * n;
      for (int64_t j = 0; j < n; ++j) row[j] += f * xr[j];
    }
    row[i] -= 1.0;
    double s = 0.0;
    for (int64_t j = 0; j < n; ++j) s += fabs(row[j]);
    if (s > mx) mx = s;
  }
  return mx;
}

}  // extern "C"
