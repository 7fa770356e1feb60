// Generator formulas shared by host code, CUDA kernels and the CPU oracle (no CUDA headers needed).
#pragma once
#include <stdint.h>

// formula ids: 1 = |i-j| (main.cpp:52), 2 = Hilbert 1/(i+j+1) (main.cpp:50),
// 3 = anti-diagonal i+1 at j = n-1-i (main.cpp:53-56), 4 = Lehmer min/max (ill-conditioned, SPD),
// 5 = diagonally dominant test matrix, 6 = |i-j| with the rows rotated by n/2 (forces off-diagonal block pivots)
#ifdef __CUDACC__
__host__ __device__
#endif
inline double jb_formula_eval(int id, int64_t i, int64_t j, int64_t n) {
  switch (id) {
    case 2: return 1.0 / (double)(i + j + 1);
    case 3: return (j == n - 1 - i) ? (double)(i + 1) : 0.0;
    case 4: { int64_t a = (i < j ? i : j) + 1, b = (i < j ? j : i) + 1; return (double)a / (double)b; }
    case 5: return (i == j) ? (double)n : 1.0 / (double)(1 + (i > j ? i - j : j - i));
    case 6: { int64_t ii = (i + n / 2) % n; return (double)(ii > j ? ii - j : j - ii); }
    default: return (double)(i > j ? i - j : j - i);
  }
}
